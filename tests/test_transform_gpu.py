"""Transform::hit in full generality on the device (transform.rs:17-165: any rotation, scale: Vec3, any child) against the
oracle's Transform wrapper over the ORIGINAL child (Cube = six rectangles with the later-side-wins tie rule of cube.rs:63-66,
rectangle uv of rectangle.rs:32,70,107, front_face against the object-space ray, Q16), plus the negative-radius sphere
(sphere.rs:44) and the BVH depth guard on a degenerate chain scene."""
import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from oracle_ffi import OracleScene
from scenes_util import compare_hits

pytestmark = pytest.mark.gpu
CAM = (((0, 0, 30), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1)


def _rays(rng, n, lo, hi, targets=None):
    o = lo + rng.random((n, 3)) * (hi - lo)
    d = rng.normal(size=(n, 3))
    if targets is not None:  # half of the rays aimed at the objects
        k = n // 2
        t = targets[rng.integers(0, len(targets), k)] + rng.normal(size=(k, 3)) * 1.5
        d[:k] = t - o[:k]
    d *= (rng.random((n, 1)) * 2 + 0.05) / np.linalg.norm(d, axis=1, keepdims=True)
    o_time = np.concatenate([o, rng.random((n, 1))], 1).astype(np.float32)
    d_tmax = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)
    seeds = rng.integers(0, 2**32, n, dtype=np.uint64).astype(np.uint32)
    return o_time, d_tmax, seeds


def _general_scene():
    white = rt.Lambertian(rt.ConstantTexture((0.7, 0.7, 0.7)))
    glass = rt.Dielectric(1.5, rt.ConstantTexture((1, 1, 1)))
    fog = rt.Isotropic(rt.ConstantTexture((0.5, 0.5, 0.9)))
    items = rt.HitableList([
        rt.Transform((20, 35, 50), (1, 1, 1), (-8, 0, 0), rt.Cube((-2, -1, -1.5), (2, 1, 1.5), white)),               # rotated Cube
        rt.Transform((0, 15, 0), (1.5, 0.5, 2.5), (0, 6, 0), rt.Cube((-1, -1, -1), (1, 1, 1), white)),                # rotated + non-uniform scale
        rt.Transform((30, 0, 10), (1, 1, 1), (8, 0, 0), rt.XZRectangle((-2, -2), (2, 2), 0.25, white)),              # rotated rectangles
        rt.Transform((0, 40, 0), (2, 1, 1), (8, 6, 0), rt.XYRectangle((-1, -1), (1, 1), 0.0, white)),
        rt.Transform((0, 0, 70), (1, 3, 1), (8, -6, 0), rt.YZRectangle((-1, -1), (1, 1), 0.0, white)),
        rt.Transform((0, 0, 0), (1, 2, 0.5), (0, 0, 0), rt.Sphere((0, 0, 0), 2, glass)),                               # ellipsoid
        rt.Transform((25, 0, 40), (3, 1, 1), (0, -6, 0), rt.Sphere((0.5, 0, 0), 1, white)),                            # rotated ellipsoid
        rt.Transform((0, 10, 0), (1, 2, 1), (-8, 6, 0), rt.MovingSphere((0, 0, 0), (1, 0.5, 0), 0.0, 1.0, 1.0, white)),
        rt.Transform((15, 25, 35), (1, 1, 1), (-8, -6, 0), rt.ConstantMedium(0.8, rt.Cube((-2, -1, -1), (2, 1, 1), white), fog)),   # rotated box medium
        rt.Transform((0, 0, 30), (2, 1, 0.7), (0, 0, 8), rt.ConstantMedium(0.6, rt.Sphere((0, 0, 0), 1.5, white), fog)),           # ellipsoidal medium
        rt.Transform((0, 0, 0), (2, 2, 2), (0, 0, -8), rt.ConstantMedium(0.4, rt.Cube((-1, -1, -1), (1, 1, 1), white), fog)),       # axis-preserving: the baked path
        rt.Sphere((0, 12, 0), 2.0, glass), rt.Sphere((0, 12, 0), -1.8, glass),                                       # hollow glass sphere
    ])
    return rt.World(items, rt.Camera(*CAM)).builder


def test_general_transforms_match_transform_hit(ctx):
    sb = _general_scene()
    d = sb.desc()
    assert d.n_prims == 13 and [d.prims[i].type for i in range(2)] == [F.PRIM_BOX, F.PRIM_BOX]
    rng = np.random.default_rng(31)
    centres = np.array([[-8, 0, 0], [0, 6, 0], [8, 0, 0], [8, 6, 0], [8, -6, 0], [0, 0, 0], [0, -6, 0], [-8, 6, 0], [-8, -6, 0], [0, 0, 8], [0, 0, -8], [0, 12, 0]], float)
    o, dd, seeds = _rays(rng, 400_000, np.array([-14.0, -11, -12]), np.array([14.0, 16, 12]), centres)
    inside = rng.normal(size=(20_000, 3))  # rays born inside the hollow sphere see its inner (negative-radius) shell first
    inside *= (1.7 * rng.random((20_000, 1)) ** (1 / 3)) / np.linalg.norm(inside, axis=1, keepdims=True)
    o[:20_000, :3] = (inside + [0, 12, 0]).astype(np.float32)
    gpu = rt.Scene(ctx, sb).intersect_batch(o, dd, seeds)
    ref_bvh = OracleScene(sb).hit_batch(o, dd, seeds)
    # The checker is the oracle's LIST arm (hit.rs:56-67, the definition of the closest hit).  Its BVH arm agrees everywhere
    # except on the negative-radius sphere: Sphere::bounding_box (sphere.rs:56-61) is center -+ radius, an INVERTED box for
    # r < 0, so upstream such a sphere is visible or not depending on the random BVH topology around it.  The device bounds
    # it by |r| and answers like the list.
    sb.set_flags(F.SCENE_FLAT_LIST)
    ref = OracleScene(sb).hit_batch(o, dd, seeds)
    sb.set_flags(0)
    differ = ref_bvh["prim"] != ref["prim"]
    print("oracle BVH arm vs list arm differ on", int(differ.sum()), "rays; list-arm primitives there:", sorted(set(ref["prim"][differ].tolist())))
    assert set(ref["prim"][differ].tolist()) <= {12}
    rep, detail = compare_hits(gpu, ref)
    print(rep, detail)
    per_prim = np.bincount(ref["prim"][ref["prim"] >= 0], minlength=13)
    print("hits per primitive", per_prim)
    bad = (gpu["prim"] != ref["prim"]) & (gpu["prim"] >= 0) & (ref["prim"] >= 0)
    print("prim confusion (gpu, ref):", sorted(set(zip(gpu["prim"][bad].tolist(), ref["prim"][bad].tolist()))))
    assert per_prim.min() > 500  # every shape is exercised
    assert rep["hit_miss"] + rep["prim"] + rep["t"] + rep["normal"] + rep["front_face"] + rep["uv"] == 0, "%s%s" % (rep, detail)
    # uv is the in-plane coordinate of the OBJECT-space rectangle (it was (0,0) when rotated shapes were lowered to triangles)
    box = (ref["prim"] == 0) | (ref["prim"] == 1)
    assert gpu["u"][box].std() > 0.2 and gpu["v"][box].std() > 0.2
    inner = ref["prim"] == 12  # negative radius: the normal points INTO the sphere, so a ray from outside sees a back face
    assert inner.sum() > 100 and np.array_equal(gpu["front_face"][inner], ref["front_face"][inner])


def test_rotated_cube_edges_and_corners(ctx):
    """Rays through the edges and corners of a rotated Cube: two or three of its six rectangles are hit at the same t and the
    later side wins (cube.rs:63-66 -> hit.rs:56-67); normal and uv identify the side."""
    white = rt.Lambertian(rt.ConstantTexture((0.7, 0.7, 0.7)))
    rot, tr = (20, 35, 50), (1, -2, 3)
    sb = rt.World(rt.HitableList([rt.Transform(rot, (1, 1, 1), tr, rt.Cube((-2, -1, -1.5), (2, 1, 1.5), white))]), rt.Camera(*CAM)).builder
    m = np.array(sb.desc().instances[0].m[:]).reshape(3, 4)
    rng = np.random.default_rng(8)
    mn, mx = np.array([-2, -1, -1.5]), np.array([2, 1, 1.5])
    os_, ds_ = [], []
    for _ in range(6000):
        c = np.where(rng.random(3) < 0.5, mn, mx)
        if rng.random() < 0.6:  # a point on an edge
            a = rng.integers(0, 3)
            c[a] = mn[a] + rng.random() * (mx[a] - mn[a])
        target = m[:, :3] @ c + m[:, 3]
        o = target + (rng.normal(size=3) * 6)
        os_.append(o); ds_.append(target - o)
    n = len(os_)
    o = np.concatenate([np.asarray(os_), np.zeros((n, 1))], 1).astype(np.float32)
    d = np.concatenate([np.asarray(ds_), np.full((n, 1), np.inf)], 1).astype(np.float32)
    gpu = rt.Scene(ctx, sb).intersect_batch(o, d, None)
    ref = OracleScene(sb).hit_batch(o, d, None)
    rep, detail = compare_hits(gpu, ref)
    print(rep)
    assert rep["hits"] > 1500
    assert rep["hit_miss"] + rep["prim"] + rep["t"] + rep["normal"] + rep["front_face"] + rep["uv"] == 0, "%s%s" % (rep, detail)


def test_images_of_general_transforms_match_oracle(ctx):
    """capture_image over the same scene: every integrator path goes through the object-space arm."""
    sb = _general_scene()
    p = rt.make_params(64, 64, 64, background=(0.6, 0.7, 0.9))
    acc, st = rt.Scene(ctx, sb).render(sb.camera, p)
    orc = OracleScene(sb)
    a, ea, _, _ = orc.render(sb.camera, rt.copy_params(p, seed=11))
    b, _, _, _ = orc.render(sb.camera, rt.copy_params(p, seed=12))
    gpu, ref = acc[..., :3].astype(np.float64) / 64, 0.5 * (a + b) / 64
    gm, rm = gpu.mean(axis=(0, 1)), ref.mean(axis=(0, 1))
    sigma = np.sqrt((((a - b) / 64) ** 2).mean(axis=(0, 1)) / 2 / (64 * 64)) * np.sqrt(1.5)
    rel = np.abs(gm - rm) / rm
    print("gpu", gm, "oracle", rm, "rel", rel, "1 sigma", sigma / rm)
    assert np.all(rel < np.maximum(0.01, 4 * sigma / rm))
    assert abs(st.extend_rays / st.paths - ea / (64 * 64 * 64)) < 0.05 * ea / (64 * 64 * 64) + 0.02


def test_chain_scene_traverses_within_the_stack(ctx):
    """The degenerate SAH input of tests/test_frontend.py on the device: World::hit equals the oracle's list scan."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    items, x, r = [], 0.0, 1e-3
    for _ in range(140):
        items.append(rt.Sphere((x + r, 0, 0), r, white))
        x += 2 * r
        r *= 1.25
    sb = rt.World(rt.HitableList(items), rt.Camera(*CAM)).builder
    st = rt.compile_stats(sb)
    assert st["stack_needed"] <= st["stack_capacity"]
    rng = np.random.default_rng(2)
    n = 100_000
    # rays from the large end towards the small one run through most of the chain
    o = np.stack([rng.random(n) * x * 1.2, rng.normal(size=n) * x * 0.02, rng.normal(size=n) * x * 0.02], 1)
    tgt = np.stack([rng.random(n) ** 8 * x, np.zeros(n), np.zeros(n)], 1) + rng.normal(size=(n, 3)) * (1e-3 + 0.02 * rng.random((n, 1)) ** 4 * x)
    d = tgt - o
    o_time = np.concatenate([o, np.zeros((n, 1))], 1).astype(np.float32)
    d_tmax = np.concatenate([d, np.full((n, 1), np.inf)], 1).astype(np.float32)
    gpu = rt.Scene(ctx, sb).intersect_batch(o_time, d_tmax, None)
    sb.set_flags(F.SCENE_FLAT_LIST)
    ref = OracleScene(sb).hit_batch(o_time, d_tmax, None)
    rep, detail = compare_hits(gpu, ref)
    print(st, rep)
    assert rep["hits"] > 20_000
    assert rep["hit_miss"] + rep["prim"] + rep["t"] + rep["normal"] + rep["front_face"] == 0, "%s%s" % (rep, detail)
