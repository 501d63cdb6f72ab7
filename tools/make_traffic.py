#!/usr/bin/env python
"""profiles/rNN/traffic_8k.json from `ncu -i X.ncu-rep --page raw --csv` dumps of the 8K step's kernels:
dram__bytes_read.sum + dram__bytes_write.sum per launch, grouped into the two passes bench.py reports
(row_pass = alternating-sum pre-kernel + row blocks; col_pass = column sums + synthetic-row pass + column blocks).
Usage: make_traffic.py out.json row.csv col.csv pre.csv"""
import csv
import json
import sys


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = []
    for d in data:
        def val(name):
            i = hdr.index(name)
            v = float(d[i].replace(",", ""))
            u = units[i].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}.get(u, 1)
        out.append({"kernel": d[hdr.index("Kernel Name")].split("(")[0], "read": val("dram__bytes_read.sum"),
                    "write": val("dram__bytes_write.sum"),
                    "us": float(d[hdr.index("gpu__time_duration.sum")].replace(",", ""))})
    return out


def main(out, row_csv, col_csv, pre_csv):
    row, col, pre = load(row_csv)[0], load(col_csv)[0], load(pre_csv)
    by = {}
    for k in pre:
        by.setdefault(k["kernel"].split("::")[-1], k)
    alt = by.get("altsum_rows_kernel")
    cs, ur = by.get("colsum_kernel"), by.get("u_rowpass_kernel")
    src = "profiles/r02/ncu_full_ols_v5_summary.txt (ncu --set full --clock-control none, one launch each)"
    t = {
        "row_pass": {"traffic": int(row["read"] + row["write"] + (alt["read"] + alt["write"] if alt else 0)), "source": src,
                     "kernels": {row["kernel"]: int(row["read"] + row["write"]),
                                 **({alt["kernel"]: int(alt["read"] + alt["write"])} if alt else {})},
                     "note": "includes the pre-kernel's re-read of the 133 MB input"},
        "col_pass": {"traffic": int(col["read"] + col["write"]), "source": src,
                     "kernels": {col["kernel"]: int(col["read"] + col["write"]),
                                 **({cs["kernel"]: int(cs["read"] + cs["write"])} if cs else {}),
                                 **({ur["kernel"]: int(ur["read"] + ur["write"])} if ur else {})},
                     "note": "traffic = the column block kernel alone (the dominant launch); its two pre-kernels run beside "
                             "the row pass in the fused step and are listed for completeness"},
    }
    json.dump(t, open(out, "w"), indent=1)
    print(json.dumps(t, indent=1))


if __name__ == "__main__":
    main(*sys.argv[1:5])
