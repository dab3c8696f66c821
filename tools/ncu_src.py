"""Summarise the source page of an .ncu-rep: per kernel, the SASS instructions with the most stall samples.

    ncu -i rep.ncu-rep --page source --csv --print-source sass > src.csv
    python tools/ncu_src.py src.csv [kernel-substring] [top-n]
"""
import csv
import sys


def main():
    path = sys.argv[1]
    want = sys.argv[2] if len(sys.argv) > 2 else ""
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
    kernels, cur, hdr = [], None, None
    for row in csv.reader(open(path)):
        if not row:
            continue
        if row[0] == "Kernel Name":
            cur = {"name": row[1], "rows": []}
            kernels.append(cur)
            hdr = None
        elif row[0] == "Address":
            hdr = row
            cur["hdr"] = hdr
        elif cur is not None and hdr is not None:
            cur["rows"].append(row)
    seen = set()
    for k in kernels:
        if want not in k["name"] or not k["rows"]:
            continue
        key = (k["name"], len(k["rows"]), k["rows"][0][0])
        if key in seen:      # ncu lists a kernel once per source view
            continue
        seen.add(key)
        hdr = k["hdr"]
        i_s, i_src, i_ex = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
        stall = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        total = sum(int(r[i_s] or 0) for r in k["rows"])
        print(f"== {k['name'][:100]}  instructions {len(k['rows'])}  samples {total}")
        agg = {}
        for r in k["rows"]:
            for i, h in stall:
                agg[h] = agg.get(h, 0) + int(r[i] or 0)
        print("   stall mix:", ", ".join(f"{h[6:]} {100 * v / max(total, 1):.1f}%" for h, v in sorted(agg.items(), key=lambda t: -t[1])[:8]))
        order = sorted(range(len(k["rows"])), key=lambda j: -int(k["rows"][j][i_s] or 0))[:top]
        for j in order:
            r = k["rows"][j]
            why = sorted(((int(r[i] or 0), h[6:]) for i, h in stall), reverse=True)[:2]
            print(f"   #{j:5d} {100 * int(r[i_s] or 0) / max(total, 1):5.1f}%  exec {r[i_ex]:>9}  {r[i_src].strip()[:70]:70s} {why}")


if __name__ == "__main__":
    main()
