"""Compact per-launch table from `ncu -i rep.ncu-rep --page raw --csv` (one row per profiled launch).

    python tools/ncu_raw_summary.py raw.csv > profiles/rNN_xxx.md
"""
import csv
import sys

COLS = [("gpu__time_duration.sum", "time"), ("sm__cycles_elapsed.avg.per_second", "SM GHz"),
        ("dram__bytes_read.sum", "DRAM rd"), ("dram__bytes_write.sum", "DRAM wr"),
        ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %"),
        ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed", "tensor pipe %"),
        ("l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem (tensor operand) %"),
        ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "smem (LSU) %"),
        ("smsp__sass_inst_executed_op_utcmma.sum", "UTCMMA instr"),
        ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
        ("lts__t_sector_hit_rate.pct", "L2 hit %"), ("launch__registers_per_thread", "regs"),
        ("launch__grid_size", "grid"), ("launch__block_size", "block")]


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    hdr, units, data = rows[0], rows[1], rows[2:]
    cols = [(hdr.index(k), n) for k, n in COLS if k in hdr]
    kn = hdr.index("Kernel Name")
    print("| kernel | " + " | ".join(f"{n} [{units[i]}]" if units[i] else n for i, n in cols) + " |")
    print("|---|" + "---|" * len(cols))
    for r in data:
        name = r[kn].split("(")[0].replace("void ", "").replace("llie::", "").replace("<unnamed>::", "").replace("unnamed>::", "")
        vals = []
        for i, _ in cols:
            try:
                v = float(r[i].replace(",", ""))
                vals.append(f"{v:.0f}" if v >= 1000 else f"{v:.3g}")
            except ValueError:
                vals.append(r[i])
        print(f"| {name} | " + " | ".join(vals) + " |")


if __name__ == "__main__":
    main()
