"""Offline: what two-ended bins would buy on a dump of tools/dbg/dump_rays.py, for several predictors of "this ray traverses for long":
the scene's dense box (b200rt_scene_bvh_quality), K boxes from median splits of the mesh inside it, an occupancy grid of it.

  python tools/dbg/analyze_split.py gpurun_out/rays_mesh_2.bin [scene=mesh]
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools", "dbg"))
import rust_raytracer_b200 as rt  # noqa: E402
from scenes_util import scene_builder  # noqa: E402
from analyze_rays import load, regroup, warp_cost  # noqa: E402

D = load(sys.argv[1])
name = sys.argv[2] if len(sys.argv) > 2 else "mesh"
sb = scene_builder(name)
q = rt.bvh_quality(sb)
print("bvh quality", q)
lo, hi = (np.array(v, np.float64) for v in q["dense_box"])
st = D["steps"]
nw = D["nw"]
dead = st == 0xffffffff
nodes = np.where(dead, 0, st & 0x3ff).astype(np.float64)
filt = np.where(dead, 0, (st >> 16) & 0xff).astype(np.float64)
exact = np.where(dead, 0, (st >> 24) & 0x3f).astype(np.float64)
leaves = filt * 0.5 + exact * 2.0
pos = (np.arange(nw)[:, None] * D["stride"] * 32 + np.arange(32)[None, :])
bin_of = np.searchsorted(D["off"], pos, side="right") - 1
bin_of = np.where(pos >= D["off"][-1], D["n_bins"], bin_of)
N = nodes.size
nf, lf, bf, deadf = nodes.ravel(), leaves.ravel(), bin_of.ravel(), dead.ravel()
of = D["rays"][:, :, 0:3].astype(np.float64).reshape(N, 3)
df = D["rays"][:, :, 4:7].astype(np.float64).reshape(N, 3)
base, eff = warp_cost(nodes, leaves)
print("as run: cost %.0f, lane efficiency %.3f; perfect cost sort %.2fx" % (base, eff, regroup(np.argsort(nf + 2 * lf, kind="stable"), nf, lf)[0] / base))


def hits_box(lo, hi, t_min=1e-3):
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / df
        t0, t1 = (lo - of) * inv, (hi - of) * inv
        tn = np.maximum(np.fmin(t0, t1).max(1), t_min)
        tf = np.fmax(t0, t1).min(1)
    return (tn <= tf) & ~deadf


def report(nm, heavy):
    c, e = regroup(np.argsort(bf * 2 + (~heavy), kind="stable"), nf, lf)
    print("%-40s heavy %5.1f %%  nodes/ray heavy %.1f light %.1f   cost %.2fx  eff %.3f" % (nm, 100 * heavy.mean(), nf[heavy].mean(), nf[~heavy & ~deadf].mean(), c / base, e))


report("dense box", hits_box(lo, hi))
# triangles of the mesh in world space
d = sb.desc()
P = np.array([[d.prims[i].p[k] for k in range(9)] for i in range(d.n_prims) if d.prims[i].type == 6 and d.prims[i].instance >= 0], np.float64).reshape(-1, 3, 3)
M = np.array([d.instances[0].m[k] for k in range(12)], np.float64).reshape(3, 4)
W = P @ M[:, :3].T + M[:, 3]
cen = W.mean(1)
print("mesh: %d triangles, world box %s .. %s" % (len(W), W.reshape(-1, 3).min(0).round(1), W.reshape(-1, 3).max(0).round(1)))


def median_boxes(idx, k):
    """k boxes from recursive median splits of the triangle centroids along the widest axis"""
    groups = [idx]
    while len(groups) < k:
        groups.sort(key=len)
        g = groups.pop()
        c = cen[g]
        ax = np.argmax(c.max(0) - c.min(0))
        o = g[np.argsort(c[:, ax], kind="stable")]
        groups += [o[:len(o) // 2], o[len(o) // 2:]]
    return [(W[g].reshape(-1, 3).min(0), W[g].reshape(-1, 3).max(0)) for g in groups]


for K in (2, 4, 8, 16, 32, 64):
    bx = median_boxes(np.arange(len(W)), K)
    h = np.zeros(N, bool)
    cnt = np.zeros(N, int)
    for blo, bhi in bx:
        hb = hits_box(blo, bhi)
        h |= hb
        cnt += hb
    report("any of %d median boxes" % K, h)
    if K >= 8:
        report("  >= 2 of %d median boxes" % K, cnt >= 2)

# occupancy grid of the dense box: a ray is heavy when it crosses an occupied voxel (sampled along the ray, 2 samples per voxel)
for G in (4, 8, 16):
    occ = np.zeros((G, G, G), bool)
    ijk = np.clip(((W.reshape(-1, 3) - lo) / (hi - lo) * G).astype(int), 0, G - 1)
    occ[ijk[:, 0], ijk[:, 1], ijk[:, 2]] = True
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / df
        t0, t1 = (lo - of) * inv, (hi - of) * inv
        tn = np.maximum(np.fmin(t0, t1).max(1), 1e-3)
        tf = np.fmax(t0, t1).min(1)
    inside = (tn <= tf) & ~deadf
    h = np.zeros(N, bool)
    S = 4 * G
    for s_ in range(S):
        t = tn + (tf - tn) * (s_ + 0.5) / S
        p = of + df * t[:, None]
        c = np.clip(((p - lo) / (hi - lo) * G), 0, G - 1e-9)
        c = np.where(np.isfinite(c), c, 0).astype(int)
        h |= inside & occ[c[:, 0], c[:, 1], c[:, 2]]
    report("occupancy grid %d^3 (%.0f %% full)" % (G, 100 * occ.mean()), h)

# ---- a finer order for the heavy class: by where the ray enters the dense box (face, G x G cell on it) instead of by bin
heavy = hits_box(lo, hi)
with np.errstate(divide="ignore", invalid="ignore"):
    inv = 1.0 / df
    t0, t1 = (lo - of) * inv, (hi - of) * inv
    tnear = np.fmin(t0, t1)
    tn = np.maximum(tnear.max(1), 1e-3)
axis = np.where(tnear.max(1) <= 1e-3, 3, tnear.argmax(1))  # 3: starts inside
pe = of + df * tn[:, None]
uvw = np.clip((pe - lo) / (hi - lo), 0, 1 - 1e-9)
uvw = np.where(np.isfinite(uvw), uvw, 0)
octant = (df[:, 0] < 0) * 1 + (df[:, 1] < 0) * 2 + (df[:, 2] < 0) * 4
for G in (2, 4, 8, 16):
    cell = (uvw * G).astype(int)
    cell_id = (cell[:, 0] * G + cell[:, 1]) * G + cell[:, 2]
    for nm, key in (("entry cell %d^3" % G, cell_id), ("entry cell %d^3, octant" % G, cell_id * 8 + octant)):
        k = np.where(heavy, 1 + key, 0).astype(np.int64)
        light_key = np.where(heavy, 0, bf).astype(np.int64)  # light rays stay in their bins
        order = np.lexsort((light_key, k))
        c, e = regroup(order, nf, lf)
        print("heavy by %-28s cost %.2fx  eff %.3f" % (nm, c / base, e))
