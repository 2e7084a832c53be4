#!/usr/bin/env python
"""Per-kernel summary of an `ncu --metrics gpu__time_duration.sum[,dram__bytes_*] --csv` launch list.
usage: launch_summary.py launches.csv [out.csv] [--seq pattern]"""
import collections
import csv
import sys


def main():
    path = sys.argv[1]
    rows = [r for r in csv.reader(open(path)) if len(r) > 10]
    hdr = rows[0]
    iN, iM, iV, iID = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("ID")
    d = collections.OrderedDict()
    for r in rows[1:]:
        d.setdefault(r[iID], {"k": r[iN]})[r[iM]] = float(r[iV].replace(",", ""))
    agg = collections.defaultdict(lambda: [0, 0.0, 0.0, 0.0])
    for v in d.values():
        a = agg[v["k"][:70]]
        a[0] += 1
        a[1] += v.get("gpu__time_duration.sum", 0)
        a[2] += v.get("dram__bytes_read.sum", 0)
        a[3] += v.get("dram__bytes_write.sum", 0)
    tot = sum(a[1] for a in agg.values())
    lines = [["kernel", "launches", "avg_us", "share", "avg_dram_read_GB", "avg_dram_write_GB"]]
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        lines.append([k, a[0], round(a[1] / a[0] / 1e3, 1), round(a[1] / tot, 4), round(a[2] / a[0] / 1e9, 3), round(a[3] / a[0] / 1e9, 3)])
    out = sys.argv[2] if len(sys.argv) > 2 and not sys.argv[2].startswith("--") else None
    if out:
        csv.writer(open(out, "w", newline="")).writerows(lines)
    for l in lines:
        print("%-72s %5s %10s %7s %8s %8s" % tuple(l))
    if "--seq" in sys.argv:
        pat = sys.argv[sys.argv.index("--seq") + 1].split("|")
        for v in d.values():
            if any(p in v["k"] for p in pat):
                print(v["k"][:40], round(v["gpu__time_duration.sum"] / 1e3, 1), round(v.get("dram__bytes_read.sum", 0) / 1e9, 2),
                      round(v.get("dram__bytes_write.sum", 0) / 1e9, 2))


if __name__ == "__main__":
    main()
