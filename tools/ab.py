"""A/B harness: bench.py (device-timed, no CPU baseline) over several builds of the library.
usage: python tools/ab.py [--steps K] name=path.so[,ENV=VAL...] ...   ("main" = the in-tree library)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
steps = "3"
args = sys.argv[1:]
if args and args[0] == "--steps":
    steps, args = args[1], args[2:]
for a in args:
    name, _, rest = a.partition("=")
    parts = rest.split(",") if rest else []
    env = dict(os.environ)
    for p in parts:
        if "=" in p:
            k, v = p.split("=", 1)
            env[k] = v
        elif p and p != "main":
            env["B200RT_LIB"] = os.path.join(ROOT, p)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", steps, "--warmup", "3", "--no-cpu-baseline", "--scaling", "weak"], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
        st = d["stage_ms_per_step"]
        print("%-14s %8.1f Mpaths/s %8.1f Mrays/s  step %7.2f ms  gen %6.2f ext %7.2f shade %6.2f  iters %5.0f  e2e %7.1f  sm %s" % (
            name, d["value"], d["mrays_per_s"], d["ms_per_step"], st["ms_generate"], st["ms_extend"], st["ms_shade"], d["wavefront_iterations_per_step"], d["e2e"]["value"], d["clocks"]["sm_mhz"]), flush=True)
    except Exception as e:  # noqa: BLE001
        print(name, "FAILED", e, r.stderr[-800:], flush=True)
