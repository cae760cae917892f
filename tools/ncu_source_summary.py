"""Summarise `ncu --page source --print-source cuda,sass --csv` per CUDA source line (samples, instructions, avg threads).
usage: ncu -i rep.ncu-rep --page source --csv --print-source cuda,sass | python tools/ncu_source_summary.py [top]"""
import csv
import sys

top = int(sys.argv[1]) if len(sys.argv) > 1 else 40
rows = list(csv.reader(sys.stdin))
cur_file, hdr, recs = None, None, []
for r in rows:
    if not r:
        continue
    if r[0] == "File Path":
        cur_file = r[1].split("/")[-1]
    elif r[0] == "Line No":
        hdr = r
    elif hdr and r[0].isdigit():
        d = dict(zip(hdr, r))
        try:
            recs.append((cur_file, int(r[0]), r[1].strip()[:110], int(d["# Samples"]), int(d["Instructions Executed"]), int(d["Thread Instructions Executed"]),
                         int(d.get("stall_long_sb", 0) or 0), int(d.get("stall_short_sb", 0) or 0), int(d.get("stall_wait", 0) or 0), int(d.get("stall_math", 0) or 0)))
        except (ValueError, KeyError):
            pass
ts, ti = sum(x[3] for x in recs), sum(x[4] for x in recs)
tt = sum(x[5] for x in recs)
print("total samples %d, warp instructions %d, thread instructions %d, avg active threads %.2f" % (ts, ti, tt, tt / max(ti, 1)))
byfile = {}
for x in recs:
    f = byfile.setdefault(x[0], [0, 0, 0])
    f[0] += x[3]; f[1] += x[4]; f[2] += x[5]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print("  %-22s samples %5.1f%%  warp-inst %5.1f%%  avg threads %.1f" % (f, 100.0 * v[0] / ts, 100.0 * v[1] / ti, v[2] / max(v[1], 1)))
print("%-20s %5s %6s %6s %5s %5s %5s %5s %5s  %s" % ("file", "line", "smp%", "inst%", "thr", "long", "short", "wait", "math", "source"))
for x in sorted(recs, key=lambda x: -x[3])[:top]:
    print("%-20s %5d %6.2f %6.2f %5.1f %5d %5d %5d %5d  %s" % (x[0], x[1], 100.0 * x[3] / ts, 100.0 * x[4] / ti, x[5] / max(x[4], 1), x[6], x[7], x[8], x[9], x[2]))
