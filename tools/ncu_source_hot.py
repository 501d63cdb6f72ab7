#!/usr/bin/env python
"""Hot spots of one kernel from `ncu -i X.ncu-rep --page source --csv` (SASS view): the instructions
that collected the most warp-stall samples, with the dominant stall reason of each.
Usage: ncu_source_hot.py src.csv [top]"""
import csv
import sys


def main(path, top=40):
    rows = list(csv.reader(open(path)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
    hdr = rows[hi]
    data = [r for r in rows[hi + 1:] if len(r) == len(hdr)]
    col = {h: i for i, h in enumerate(hdr)}
    stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
    tot = sum(int(r[col["# Samples"]] or 0) for r in data)
    print("kernel:", rows[0][1][:150] if len(rows[0]) > 1 else "")
    print("instructions:", len(data), "samples:", tot)
    agg = {}
    for h in stall_cols:
        agg[h] = sum(int(r[col[h]] or 0) for r in data)
    print("stall totals:", {k: v for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v})
    order = sorted(range(len(data)), key=lambda i: -int(data[i][col["# Samples"]] or 0))[:top]
    for i in sorted(order):
        r = data[i]
        st = sorted(((int(r[col[h]] or 0), h[6:]) for h in stall_cols), reverse=True)[:2]
        print(f"{i:5d} {int(r[col['# Samples']]):6d}  {r[col['Source']][:70]:70s} {st}")


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 40)
