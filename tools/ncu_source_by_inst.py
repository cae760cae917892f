"""Per-source-line warp-instruction counts (aggregated over inlined copies). usage: ncu -i rep --page source --csv --print-source cuda,sass | python tools/ncu_source_by_inst.py [top]"""
import csv, sys, collections
top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
want = sys.argv[2] if len(sys.argv) > 2 else None  # optional kernel-name substring filter
keep = True
rows = list(csv.reader(sys.stdin))
cur, hdr = None, None
agg = collections.defaultdict(lambda: [0, 0, 0, ""])
for r in rows:
    if not r: continue
    if r[0] == "Kernel Name": keep = (want is None) or (want in r[1])
    if not keep: continue
    if r[0] == "File Path": cur = r[1].split("/")[-1]
    elif r[0] == "Line No": hdr = r
    elif hdr and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            a = agg[(cur, int(r[0]))]
            a[0] += int(d["Instructions Executed"]); a[1] += int(d["Thread Instructions Executed"]); a[2] += int(d["# Samples"]); a[3] = r[1].strip()[:100]
        except (ValueError, KeyError): pass
ti = sum(a[0] for a in agg.values()); ts = sum(a[2] for a in agg.values())
print("total warp-inst %d samples %d" % (ti, ts))
for (f, l), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print("%-18s %4d inst %5.2f%% (%9d) thr %4.1f smp %5.2f%%  %s" % (f, l, 100.0 * a[0] / ti, a[0], a[1] / max(a[0], 1), 100.0 * a[2] / ts, a[3]))
