"""profiles/ncu_traffic.json from `ncu -i rep.ncu-rep --page raw --csv` dumps of one `--set full` capture per workload.

    python tools/ncu_traffic.py --dce raw_dce.csv 16 --msrcr raw_msrcr.csv 8 [--msrcr-u8 raw.csv 8] --capture "r02 ..."

Writes, per kernel, dram__bytes_read.sum + dram__bytes_write.sum per PIXEL of the workload (summed over that
kernel's launches in one call), the whole-step figure, and the hash of the kernel sources of the build that was
profiled (bench.py compares it with its own and reports `traffic_source.matches_this_build`).
"""
import argparse
import csv
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

UNIT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
CHAIN = {"0": "dce_chain_c123_kernel", "1": "dce_chain_c45_kernel", "2": "dce_chain_c67_kernel",
         "7": "dce_split_first_kernel", "8": "dce_split_k32_kernel", "9": "dce_split_k64_kernel", "10": "dce_split_last_kernel"}


def bench_name(ncu_name: str) -> str:
    m = re.search(r"dce_chain_kernel<\(?int\)?(\d+)>|dce_chain_kernel<(\d+)>", ncu_name)
    if m:
        gid = m.group(1) or m.group(2)
        return CHAIN.get(gid, f"dce_chain_kernel<{gid}>")
    m = re.search(r"([a-z0-9_]+_kernel)", ncu_name)
    base = m.group(1) if m else ncu_name
    return {"fir_x8_kernel": "fir_x_kernel", "fir_y8_kernel": "fir_y_kernel"}.get(base, base)


def per_kernel(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    kn, rd, wr = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
    out = {}
    for r in data:
        b = float(r[rd].replace(",", "")) * UNIT[units[rd]] + float(r[wr].replace(",", "")) * UNIT[units[wr]]
        out[bench_name(r[kn])] = out.get(bench_name(r[kn]), 0.0) + b
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dce", nargs=2, metavar=("CSV", "FRAMES"))
    ap.add_argument("--msrcr", nargs=2, metavar=("CSV", "FRAMES"))
    ap.add_argument("--msrcr-u8", nargs=2, metavar=("CSV", "FRAMES"))
    ap.add_argument("--dce32", nargs=2, metavar=("CSV", "FRAMES"), help="Zero-DCE fp32 (hi/lo split) on 1080p frames")
    ap.add_argument("--capture", default="")
    a = ap.parse_args()
    import bench
    doc = {"source_hash": bench.csrc_hash(), "capture": a.capture, "kernels": {}}
    if a.dce:
        px = int(a.dce[1]) * 1920 * 1080
        k = per_kernel(a.dce[0])
        doc["kernels"].update({n: b / px for n, b in k.items()})
        doc["dce_step_bytes_per_px"] = sum(k.values()) / px
    if a.dce32:
        px = int(a.dce32[1]) * 1920 * 1080
        k = per_kernel(a.dce32[0])
        k = {("dce_pack_input_split_kernel" if n == "dce_pack_input_kernel" else n): b for n, b in k.items()}
        doc["kernels"].update({n: b / px for n, b in k.items()})
        doc["dce32_step_bytes_per_px"] = sum(k.values()) / px
    if a.msrcr:
        px = int(a.msrcr[1]) * 3840 * 2160
        k = per_kernel(a.msrcr[0])
        doc["kernels"].update({n: b / px for n, b in k.items()})
        doc["msrcr_step_bytes_per_px"] = sum(k.values()) / px
    if a.msrcr_u8:
        px = int(a.msrcr_u8[1]) * 3840 * 2160
        k = per_kernel(a.msrcr_u8[0])
        doc["kernels"].update({n + ":u8": b / px for n, b in k.items()})
        doc["msrcr_step_bytes_per_px_u8"] = sum(k.values()) / px
    json.dump(doc, open(os.path.join(ROOT, "profiles", "ncu_traffic.json"), "w"), indent=1, sort_keys=True)
    print(json.dumps(doc, indent=1, sort_keys=True))


if __name__ == "__main__":
    main()
