"""Debug helper: the degenerate chain scene on the device, BVH arm vs flat-list arm vs the oracle's list scan."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from oracle_ffi import OracleScene

ratio = float(sys.argv[1]) if len(sys.argv) > 1 else 1.25
count = int(sys.argv[2]) if len(sys.argv) > 2 else 140
white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
items, x, r = [], 0.0, 1e-3
radii = []
for _ in range(count):
    items.append(rt.Sphere((x + r, 0, 0), r, white)); radii.append(r)
    x += 2 * r; r *= ratio
sb = rt.World(rt.HitableList(items), rt.Camera(((0, 0, 30), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1)).builder
print(rt.compile_stats(sb), "chain length %.3g" % x)
rng = np.random.default_rng(2)
n = 100_000
o = np.stack([rng.random(n) * x * 1.2, rng.normal(size=n) * x * 0.02, rng.normal(size=n) * x * 0.02], 1)
tgt = np.stack([rng.random(n) ** 8 * x, np.zeros(n), np.zeros(n)], 1) + rng.normal(size=(n, 3)) * (1e-3 + 0.02 * rng.random((n, 1)) ** 4 * x)
d = tgt - o
o_time = np.concatenate([o, np.zeros((n, 1))], 1).astype(np.float32)
d_tmax = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)
ctx = rt.Context(0)
g_bvh = rt.Scene(ctx, sb).intersect_batch(o_time, d_tmax, None)
sb.set_flags(F.SCENE_FLAT_LIST)
g_flat = rt.Scene(ctx, sb).intersect_batch(o_time, d_tmax, None)
ref = OracleScene(sb).hit_batch(o_time, d_tmax, None)
for name, g in (("bvh", g_bvh), ("flat", g_flat)):
    hm = (g["prim"] >= 0) != (ref["prim"] >= 0)
    both = (g["prim"] >= 0) & (ref["prim"] >= 0)
    pd = both & (g["prim"] != ref["prim"])
    print(name, "hit/miss diffs", int(hm.sum()), "prim diffs", int(pd.sum()), "of", n)
    bad = np.flatnonzero(hm | pd)[:8]
    for i in bad:
        print("  ray", i, "o", o_time[i, :3], "d", d_tmax[i, :3], "gpu prim", g["prim"][i], "t", g["t"][i], "| ref prim", ref["prim"][i], "t %.6g" % ref["t"][i], "r_ref %.3g" % radii[ref["prim"][i]] if ref["prim"][i] >= 0 else "")
    if (hm | pd).any():
        rp = ref["prim"][hm | pd]
        print("  missed reference prims histogram:", np.bincount(rp[rp >= 0], minlength=count)[-40:])
