"""Per-kernel totals from an `ncu --metrics gpu__time_duration.sum --csv` launch list. usage: python tools/ncu_launch_summary.py file.csv"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
agg = collections.defaultdict(lambda: [0, 0.0])
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    if r[ui] == "ns":
        v /= 1e3
    k = r[ki].split("(")[0]
    agg[k][0] += 1
    agg[k][1] += v
tot = sum(v[1] for v in agg.values())
print("%-44s %6s %12s %7s %9s" % ("kernel", "n", "total_us", "share", "avg_us"))
for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
    print("%-44s %6d %12.1f %7.3f %9.1f" % (k[:44], v[0], v[1], v[1] / tot, v[1] / v[0]))
