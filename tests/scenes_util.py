"""Shared scene / ray-batch helpers for the parity tests."""
import os

import numpy as np

import rust_raytracer_b200 as rt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
DATA = os.path.join(ROOT, "data")


_EARTH = None


def earth_rgb():
    """Decoded data/earthmap.jpg, 1024x512 — the texture bench.py renders with (decode once on the host: the same RGB8
    array feeds oracle and GPU, so the decoder choice cancels out); the 256x128 fixture only if PIL is missing."""
    global _EARTH
    if _EARTH is None:
        try:
            from PIL import Image
            _EARTH = np.ascontiguousarray(np.asarray(Image.open(os.path.join(DATA, "earthmap.jpg")).convert("RGB"), dtype=np.uint8))
        except Exception:
            _EARTH = np.load(os.path.join(ROOT, "tests", "golden", "earthmap_256x128.npy"))
    return _EARTH


AUX = {"cornell_ref": os.path.join(DATA, "mesh", "cube.obj"), "mesh": os.path.join(DATA, "mesh", "bun315.obj"),
       "mesh_small": os.path.join(DATA, "mesh", "bun315.obj")}


def oracle_scene(name, seed=1, tri_pad=0.0):
    """The oracle's OWN construction of a canonical scene / JSON file (oracle/oracle_scenes.inc): built inside liboracle.so as
    the reference's object graph, never from the product's tables."""
    from oracle_ffi import OracleScene
    if name.startswith("json:"):
        return OracleScene.load_json(open(os.path.join(DATA, name[5:] + ".json")).read())
    return OracleScene.make(name, seed=seed, aux_path=AUX.get(name), earth=earth_rgb() if name == "book2" else None, tri_pad=tri_pad)


def scene_builder(name, seed=1):
    if name.startswith("json:"):
        path = os.path.join(DATA, name[5:] + ".json")
        sb = rt.SceneBuilder().load_json(open(path).read())
        sb.params = rt.make_params(192, 108, 8)
        return sb
    aux = AUX.get(name)
    earth = earth_rgb() if name == "book2" else None
    return rt.SceneBuilder().make_scene(name, seed=seed, aux_path=aux, earth=earth)


def camera_rays(cam, width, height, rng):
    """Primary rays through pixel centres with jitter (camera.rs:97-99, lens offset ignored)."""
    ys, xs = np.mgrid[0:height, 0:width]
    s = (xs + rng.random(xs.shape)) / (width - 1)
    t = 1.0 - (ys + rng.random(ys.shape)) / (height - 1)
    o = np.array(cam.origin[:], np.float64)
    llc, hor, ver = (np.array(v[:], np.float64) for v in (cam.lower_left_corner, cam.horizontal, cam.vertical))
    d = llc + s[..., None] * hor + t[..., None] * ver - o
    n = width * height
    return np.broadcast_to(o, (n, 3)).astype(np.float32), d.reshape(n, 3).astype(np.float32)


def scene_bounds(desc):
    lo, hi = np.full(3, np.inf), np.full(3, -np.inf)
    for i in range(desc.n_prims):
        p = desc.prims[i]
        v = np.array(p.p[:], np.float64)
        if p.instance >= 0:
            continue
        if p.type in (0, 1, 7):
            r = min(v[3], 2000.0)
            lo, hi = np.minimum(lo, v[:3] - r), np.maximum(hi, v[:3] + r)
        elif p.type in (5, 8):
            lo, hi = np.minimum(lo, v[:3]), np.maximum(hi, v[3:6])
        elif p.type == 6:
            for k in range(3):
                lo, hi = np.minimum(lo, v[3 * k:3 * k + 3]), np.maximum(hi, v[3 * k:3 * k + 3])
    if not np.all(np.isfinite(lo)):
        lo, hi = np.zeros(3), np.full(3, 555.0)
    return lo, hi


def ray_batch(sb, n_random, rng, cam_res=(96, 64)):
    """Camera rays + uniformly random rays inside the scene bounds + secondary rays leaving hit surfaces."""
    desc = sb.desc()
    o1, d1 = camera_rays(sb.camera, cam_res[0], cam_res[1], rng)
    lo, hi = scene_bounds(desc)
    ext = hi - lo
    o2 = (lo - 0.1 * ext + rng.random((n_random, 3)) * 1.2 * ext).astype(np.float32)
    d2 = rng.normal(size=(n_random, 3))
    d2 *= (rng.random((n_random, 1)) * 2 + 0.05) / np.linalg.norm(d2, axis=1, keepdims=True)  # unnormalised, like the reference
    o = np.concatenate([o1, o2]).astype(np.float32)
    d = np.concatenate([d1, d2.astype(np.float32)]).astype(np.float32)
    n = o.shape[0]
    time = rng.random(n).astype(np.float32)
    o_time = np.concatenate([o, time[:, None]], axis=1).astype(np.float32)
    d_tmax = np.concatenate([d, np.full((n, 1), np.inf, np.float32)], axis=1).astype(np.float32)
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    return o_time, d_tmax, seeds


def compare_hits(gpu, ref, rtol=1e-5, label=""):
    """hit/miss and primitive identical (exact ties excepted), t and normal within rtol (north_star)."""
    g_hit, r_hit = gpu["prim"] >= 0, ref["prim"] >= 0
    both = g_hit & r_hit
    dt = np.zeros(len(gpu))
    dt[both] = np.abs(gpu["t"][both].astype(np.float64) - ref["t"][both]) / np.maximum(np.abs(ref["t"][both]), 1e-30)
    hm = np.flatnonzero(g_hit != r_hit)
    prim_diff = both & (gpu["prim"] != ref["prim"])
    # a different primitive at the same t is a tie (two surfaces meeting along an edge): bvh.rs:91-101 resolves it
    # by traversal order, which is random upstream
    tie = prim_diff & (dt <= 1e-6)
    bad_prim = np.flatnonzero(prim_diff & ~tie)
    same = both & ~prim_diff
    bad_t = np.flatnonzero(same & (dt > rtol))
    gn = np.stack([gpu["nx"], gpu["ny"], gpu["nz"]], 1).astype(np.float64)
    rn = np.stack([ref["nx"], ref["ny"], ref["nz"]], 1)
    dn = np.abs(gn - rn).max(axis=1)
    bad_n = np.flatnonzero(same & (dn > rtol))
    bad_ff = np.flatnonzero(same & (gpu["front_face"] != ref["front_face"]))
    duv = np.maximum(np.abs(gpu["u"] - ref["u"]), np.abs(gpu["v"] - ref["v"]))
    duv = np.minimum(duv, np.abs(1.0 - duv))  # the sphere u seam wraps
    bad_uv = np.flatnonzero(same & (duv > 1e-4))
    report = dict(n=len(gpu), hits=int(r_hit.sum()), hit_miss=len(hm), prim=len(bad_prim), ties=int(tie.sum()), t=len(bad_t), normal=len(bad_n),
                  front_face=len(bad_ff), uv=len(bad_uv), max_dt=float(dt[same].max()) if same.any() else 0.0, max_dn=float(dn[same].max()) if same.any() else 0.0)
    detail = ""
    for name, idx in (("hit_miss", hm), ("prim", bad_prim), ("t", bad_t), ("normal", bad_n), ("front_face", bad_ff), ("uv", bad_uv)):
        for i in idx[:3]:
            detail += "\n  %s[%d]: gpu=%s ref=(t=%.9g prim=%d n=(%.6f %.6f %.6f) ff=%d uv=(%.6f,%.6f))" % (
                name, i, gpu[i], ref["t"][i], ref["prim"][i], ref["nx"][i], ref["ny"][i], ref["nz"][i], ref["front_face"][i], ref["u"][i], ref["v"][i])
    return report, detail
