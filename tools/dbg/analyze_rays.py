"""Offline analysis of a tools/dbg/dump_rays.py file: lane occupancy of the BVH traversal in k_wf_extend, where the
idle lanes are, and what alternative queue orders would buy.

Cost model of one 32-ray batch: max over lanes of (inner nodes) * C_NODE + max over lanes of (leaves) * C_LEAF
(a warp runs its node / leaf phases as long as its slowest lane; while-while traversal).  The lane-weighted
efficiency is sum(work) / (32 * max).

  python tools/dbg/analyze_rays.py gpurun_out/rays_book2_12.bin
"""
import sys

import numpy as np

C_NODE, C_LEAF = 1.0, 2.0


def load(path):
    raw = np.fromfile(path, dtype=np.uint32)
    n_ext, stride, nw, n_bins, bin_mul, n_prims, n_bvh, _ = (int(x) for x in raw[:8])
    at = 8
    off = raw[at:at + n_bins + 1].astype(np.int64); at += n_bins + 1
    root = raw[at:at + 32].view(np.float32).reshape(8, 4); at += 32
    rays = raw[at:at + nw * 256].view(np.float32).reshape(nw, 32, 8); at += nw * 256
    hits = raw[at:at + nw * 128].reshape(nw, 32, 4); at += nw * 128
    steps = raw[at:at + nw * 32].reshape(nw, 32); at += nw * 32
    return dict(n_ext=n_ext, stride=stride, nw=nw, n_bins=n_bins, bin_mul=bin_mul, n_prims=n_prims, n_bvh=n_bvh, off=off, root=root, rays=rays, hits=hits, steps=steps)


def warp_cost(nodes, leaves):
    """nodes, leaves: (n_warps, 32) -> total cost, lane efficiency"""
    c = nodes.max(1) * C_NODE + leaves.max(1) * C_LEAF
    work = (nodes * C_NODE + leaves * C_LEAF).sum()
    return c.sum(), work / (32.0 * c.sum())


def regroup(order, nodes, leaves):
    n = (len(order) // 32) * 32
    o = order[:n]
    return warp_cost(nodes[o].reshape(-1, 32), leaves[o].reshape(-1, 32))


def main(path):
    D = load(path)
    nw = D["nw"]
    st = D["steps"]
    dead = st == 0xffffffff
    nodes = np.where(dead, 0, st & 0x3ff).astype(np.float64)
    leaves = np.where(dead, 0, (st >> 10) & 0x3f).astype(np.float64)
    filt = np.where(dead, 0, (st >> 16) & 0xff).astype(np.float64)
    exact = np.where(dead, 0, (st >> 24) & 0x3f).astype(np.float64)
    print("per ray: f32 filter calls %.2f (batch max %.2f), f64 decisions %.2f (batch max %.2f)" % (filt.mean(), filt.max(1).mean(), exact.mean(), exact.max(1).mean()))
    leaves = leaves + 0.0  # leaf phase cost proxy below: filters and exact tests weigh more than the leaf visit itself
    leaves = filt * 0.5 + exact * 2.0
    pos = (np.arange(nw)[:, None] * D["stride"] * 32 + np.arange(32)[None, :])
    bin_of = np.searchsorted(D["off"], pos, side="right") - 1  # bin of each queue position (new paths sit behind the last bin)
    bin_of = np.where(pos >= D["off"][-1], D["n_bins"], bin_of)
    o = D["rays"][:, :, 0:3].astype(np.float64)
    d = D["rays"][:, :, 4:7].astype(np.float64)
    print("n_ext %d, sampled batches %d (every %d-th), dead %.2f %%, new paths %.1f %%" % (D["n_ext"], nw, D["stride"], 100 * dead.mean(), 100 * (bin_of == D["n_bins"]).mean()))
    print("per ray: inner nodes %.2f, leaves %.2f" % (nodes.mean(), leaves.mean()))
    base, eff = warp_cost(nodes, leaves)
    print("as run:            cost %10.0f  lane efficiency %.3f (nodes only: %.3f)" % (base, eff, nodes.sum() / (32 * nodes.max(1).sum())))
    # where the waste is: by bin class
    fog_bin = 1 + ((D["n_bvh"] * D["bin_mul"]) >> 32)  # first out-of-BVH primitive
    cls = np.where(bin_of == D["n_bins"], 2, np.where(bin_of >= fog_bin, 1, 0))
    wcls = np.array([np.bincount(c, minlength=3).argmax() for c in cls])
    for k, nm in enumerate(["in-BVH prim bins", "out-of-BVH (fog) bins", "new camera paths"]):
        m = wcls == k
        if m.any():
            c, e = warp_cost(nodes[m], leaves[m])
            print("  %-22s batches %5.1f %%  cost share %5.1f %%  efficiency %.3f  nodes/ray %.2f  max nodes/batch %.1f" % (nm, 100 * m.mean(), 100 * c / base, e, nodes[m].mean(), nodes[m].max(1).mean()))
    # histogram of node counts
    h = np.bincount(nodes.astype(int).ravel(), minlength=40)
    print("node-count histogram (share of rays): " + " ".join("%d:%.1f" % (i, 100 * h[i] / h.sum()) for i in range(min(len(h), 32)) if h[i] / h.sum() > 0.005))
    # ---- alternative orders (population = the sampled rays, regrouped into batches of 32)
    N = nodes.size
    nf, lf, bf = nodes.ravel(), leaves.ravel(), bin_of.ravel()
    of, df = o.reshape(N, 3), d.reshape(N, 3)
    ident = np.arange(N)
    print("regroup identity:  cost %10.0f" % regroup(ident, nf, lf)[0])
    rng = np.random.default_rng(0)
    c, e = regroup(rng.permutation(N), nf, lf)
    print("random order:      cost %10.0f (%.2fx)  eff %.3f" % (c, c / base, e))
    c, e = regroup(np.argsort(nf * C_NODE + lf * C_LEAF, kind="stable"), nf, lf)
    print("perfect cost sort: cost %10.0f (%.2fx)  eff %.3f   <- lower bound of any reorder" % (c, c / base, e))
    # root box test of each ray (slab test against the union of the root's children)
    r = D["root"]
    lo = np.array([r[0].min(), r[2].min(), r[4].min()], np.float64)
    hi = np.array([r[1][np.isfinite(r[1])].max(), r[3][np.isfinite(r[3])].max(), r[5][np.isfinite(r[5])].max()], np.float64)
    lo = np.array([r[0][np.isfinite(r[0])].min(), r[2][np.isfinite(r[2])].min(), r[4][np.isfinite(r[4])].min()], np.float64)
    with np.errstate(divide="ignore", invalid="ignore"):
        inv = 1.0 / df
        t0, t1 = (lo - of) * inv, (hi - of) * inv
        tn = np.minimum(t0, t1).max(1)
        tf = np.maximum(t0, t1).min(1)
    root_hit = (tn <= tf) & (tf >= 0.001)
    span = np.where(root_hit, tf - np.maximum(tn, 0.001), 0.0) * np.linalg.norm(df, axis=1)
    print("root box %s .. %s; rays hitting it: %.1f %%; nodes/ray if hit %.2f, if miss %.2f" % (lo.round(1), hi.round(1), 100 * root_hit.mean(), nf[root_hit].mean(), nf[~root_hit].mean() if (~root_hit).any() else 0))
    octant = (df[:, 0] < 0) * 1 + (df[:, 1] < 0) * 2 + (df[:, 2] < 0) * 4
    keys = {
        "bin, root hit":           bf * 2 + root_hit,
        "bin, octant":             bf * 8 + octant,
        "bin, root hit, octant":   (bf * 2 + root_hit) * 8 + octant,
        "bin, span quartile":      bf * 4 + np.digitize(span, np.quantile(span[root_hit], [0.33, 0.66]) if root_hit.any() else [1, 2]) * root_hit,
        "root hit only (no bin)":  root_hit * 1,
    }
    for nm, k in keys.items():
        c, e = regroup(np.argsort(k, kind="stable"), nf, lf)
        print("key %-24s cost %10.0f (%.2fx)  eff %.3f" % (nm, c, c / base, e))
    # warp-level two-batch repack: a warp takes 64 consecutive rays, runs the root-hitters first (packed), then the rest
    n64 = (N // 64) * 64
    tot = 0.0
    for w in range(0, n64, 64):
        idx = np.arange(w, w + 64)
        k = np.argsort(~root_hit[idx], kind="stable")
        a, b = idx[k[:32]], idx[k[32:]]
        tot += nf[a].max() * C_NODE + lf[a].max() * C_LEAF + nf[b].max() * C_NODE + lf[b].max() * C_LEAF
    print("warp repack of 64 by root hit: cost %10.0f (%.2fx)" % (tot, tot / regroup(ident[:n64], nf, lf)[0]))
    for G in (128, 256, 1024):
        nG = (N // G) * G
        tot = 0.0
        for w in range(0, nG, G):
            idx = np.arange(w, w + G)
            k = idx[np.argsort(nf[idx] * C_NODE + lf[idx] * C_LEAF, kind="stable")]
            tot += (nf[k].reshape(-1, 32).max(1) * C_NODE + lf[k].reshape(-1, 32).max(1) * C_LEAF).sum()
        print("perfect sort inside groups of %4d: cost %10.0f (%.2fx)" % (G, tot, tot / regroup(ident[:nG], nf, lf)[0]))


if __name__ == "__main__":
    main(sys.argv[1])
