"""Dynamic opcode histogram from `ncu -i rep --page source --csv --print-source sass` (first kernel in the report).
usage: ncu -i rep.ncu-rep --page source --csv --print-source sass | python tools/ncu_sass_opcodes.py [top]"""
import collections
import csv
import sys

top = int(sys.argv[1]) if len(sys.argv) > 1 else 30
rows = list(csv.reader(sys.stdin))
hdr = None
agg = collections.defaultdict(lambda: [0, 0, 0])
kernels = 0
for r in rows:
    if not r:
        continue
    if r[0] == "Kernel Name":
        kernels += 1
        if kernels > 1:
            break
    elif r[0] == "Address":
        hdr = r
    elif hdr and r[0].startswith("0x"):
        d = dict(zip(hdr, r))
        toks = d["Source"].split()
        op = toks[1] if toks and toks[0].startswith("@") else toks[0]
        op = op.split(".")[0]
        a = agg[op]
        a[0] += int(d["Instructions Executed"]); a[1] += int(d["Thread Instructions Executed"]); a[2] += int(d["# Samples"])
ti = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("warp instructions %d, samples %d" % (ti, ts))
for op, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-10s inst %6.2f%%  thr/inst %5.1f  samples %6.2f%%" % (op, 100.0 * a[0] / ti, a[1] / max(a[0], 1), 100.0 * a[2] / max(ts, 1)))
