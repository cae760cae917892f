"""Oracle pinned against what the reference ships: the bounding boxes stored in data/scene_*.json (a known answer
for Sphere::bounding_box + AABB::surrounding_box, sphere.rs:56-61, aabb.rs:33-45) and hand-derived analytic cases
for hit / scatter / the render loop (the reference has no tests for those: parity unpinned)."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from oracle_ffi import OracleScene, lib
from scenes_util import DATA, scene_builder


def _walk(node, out):
    t = node["type"]
    if t == "Sphere":
        out.append(node)
    elif t == "HitableList":
        for it in node["items"]:
            _walk(it, out)
    elif t == "BVHNode":
        _walk(node["left"], out)
        _walk(node["right"], out)


def _bvh_roots(node, out):
    if node["type"] == "BVHNode":
        out.append(node)
    elif node["type"] == "HitableList":
        for it in node["items"]:
            _bvh_roots(it, out)


@pytest.mark.parametrize("name", ["scene_10", "scene_500"])
def test_json_bounding_boxes_are_reproduced(name):
    doc = json.load(open(os.path.join(DATA, name + ".json")))
    roots = []
    _bvh_roots(doc["objects"], roots)
    assert roots
    for root in roots:
        spheres = []
        _walk(root, spheres)
        prims = (F.Prim * len(spheres))()
        for i, s in enumerate(spheres):
            prims[i].type = F.PRIM_SPHERE
            c = s["center"]
            # the stored boxes were computed in f64 from the f64 literals: feed the oracle f64-exact values where
            # f32 can hold them, and compare with the tolerance of the f32 rounding of the inputs
            prims[i].p[0], prims[i].p[1], prims[i].p[2], prims[i].p[3] = c["x"], c["y"], c["z"], s["radius"]
        out = np.zeros(6)
        lib().oracle_sphere_union_box(prims, len(spheres), out.ctypes.data_as(C.c_void_p))
        bb = root["bounding_box"]
        want = np.array([bb["min"]["x"], bb["min"]["y"], bb["min"]["z"], bb["max"]["x"], bb["max"]["y"], bb["max"]["z"]])
        np.testing.assert_allclose(out, want, rtol=0, atol=1e-6)  # f32 rounding of centre/radius at |x| ~ 5


def _one_sphere(material, center=(0, 0, -1), radius=0.5):
    w = rt.World(rt.HitableList([rt.Sphere(center, radius, material)]), rt.Camera(((0, 0, 0), (0, 0, -1)), (0, 1, 0), 90, 1.0, 0.0, 1.0))
    return w.builder


def test_sphere_hit_closed_form():
    sb = _one_sphere(rt.Lambertian(rt.ConstantTexture((0.5, 0.5, 0.5))))
    orc = OracleScene(sb)
    o = np.array([[0, 0, 0, 0], [0, 0, 0, 0], [0, 0, -1, 0], [0, 2, 0, 0]], np.float32)
    d = np.array([[0, 0, -1, np.inf], [0, 0, -2, np.inf], [0, 0, -1, np.inf], [0, 0, -1, np.inf]], np.float32)
    h = orc.hit_batch(o, d)
    assert h["t"][0] == 0.5 and h["front_face"][0] == 1 and (h["nx"][0], h["ny"][0], h["nz"][0]) == (0.0, 0.0, 1.0)
    assert h["t"][1] == 0.25  # unnormalised direction: t scales (ray.rs:12-14)
    assert h["t"][2] == 0.5 and h["front_face"][2] == 0 and h["nz"][2] == 1.0  # from the centre: exit root, normal flipped
    assert h["prim"][3] == -1
    # get_uv (sphere.rs:16-20) at the +z pole of the outward normal (0,0,1): phi = atan2(-1, 0) + pi = pi/2, theta = pi/2
    assert abs(h["u"][0] - 0.25) < 1e-15 and abs(h["v"][0] - 0.5) < 1e-15


def test_rect_cube_triangle_transform_closed_form():
    white = rt.Lambertian(rt.ConstantTexture((0.7, 0.7, 0.7)))
    tri = rt.Triangle((0, 0, 0), (1, 0, 0), (0, 1, 0), (0, 0, 1), (0, 0, 1), (0, 0, 1), white)
    items = rt.HitableList([
        rt.XYRectangle((0, 0), (2, 4), -3.0, white),                      # prim 0
        rt.Cube((10, 10, 10), (12, 12, 12), white),                        # prim 1
        rt.Transform((0, 0, 0), (2, 2, 2), (20, 0, 0), tri),               # prim 2: scaled x2, moved to x = 20
        rt.Transform((0, 90, 0), (1, 1, 1), (0, 0, 50), rt.Sphere((1, 0, 0), 0.5, white)),  # prim 3: rotated about Y
    ])
    sb = rt.World(items, rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1.0, 0.0, 1.0)).builder
    orc = OracleScene(sb)
    o = np.array([[1, 1, 0, 0], [11, 11, 0, 0], [11, 11, 11, 0], [20.5, 0.5, 5, 0], [0, 0, 40, 0]], np.float32)
    d = np.array([[0, 0, -1, np.inf], [0, 0, 1, np.inf], [1, 0, 0, np.inf], [0, 0, -1, np.inf], [0, 0, 1, np.inf]], np.float32)
    h = orc.hit_batch(o, d)
    assert h["prim"][0] == 0 and h["t"][0] == 3.0 and h["nz"][0] == 1.0 and h["front_face"][0] == 1 and (h["u"][0], h["v"][0]) == (0.5, 0.25)
    assert h["prim"][1] == 1 and h["t"][1] == 10.0 and h["nz"][1] == -1.0        # enters through the z = 10 face
    assert h["prim"][2] == 1 and h["t"][2] == 1.0 and h["nx"][2] == -1.0         # from inside: hits the exit face (Q18)
    assert h["prim"][3] == 2 and h["t"][3] == 5.0 and h["nz"][3] == 1.0 and h["front_face"][3] == 1
    # RotateY(90): x -> -z, so the sphere centre (1,0,0) lands at (0,0,-1) + (0,0,50) = (0,0,49); hit at z = 48.5.
    # SURVEY Q16 (transform.rs:151-165): set_face_normal tests the OBJECT-space direction d' = (-1,0,0) against the
    # WORLD-space normal M*n' = (0,0,-1): d'.n = 0 is not < 0, so front_face = false and the normal is flipped to
    # (0,0,+1) — along the ray.  This is the reference's behaviour for rotating Transforms and the oracle keeps it.
    assert h["prim"][4] == 3 and abs(h["t"][4] - 8.5) < 1e-12 and abs(h["nz"][4] - 1.0) < 1e-12 and h["front_face"][4] == 0


def test_constant_medium_uses_the_injected_draw():
    iso = rt.Isotropic(rt.ConstantTexture((1, 1, 1)))
    items = rt.HitableList([rt.ConstantMedium(0.5, rt.Sphere((0, 0, -5), 2.0, iso), iso)])
    sb = rt.World(items, rt.Camera(((0, 0, 0), (0, 0, -1)), (0, 1, 0), 40, 1.0, 0.0, 1.0)).builder
    orc = OracleScene(sb)
    n = 20000
    o = np.zeros((n, 4), np.float32)
    d = np.tile(np.array([0, 0, -1, np.inf], np.float32), (n, 1))
    seeds = np.arange(n, dtype=np.uint32)
    h = orc.hit_batch(o, d, seeds)
    hit = h["prim"] >= 0
    # chord length 4 at density 0.5: P(hit) = 1 - exp(-2); hits lie in [3, 7] and follow the exponential law
    assert abs(hit.mean() - (1 - np.exp(-2.0))) < 0.01
    assert h["t"][hit].min() >= 3.0 and h["t"][hit].max() <= 7.0
    assert abs(np.mean(h["t"][hit] - 3.0 < 1.0) - (1 - np.exp(-0.5)) / (1 - np.exp(-2.0))) < 0.015
    # outward normal (1,0,0) (medium.rs:47); d.n = 0 is not < 0 -> front_face false and the normal is negated (hit.rs:25-30)
    assert np.all(h["nx"][hit] == -1.0) and np.all(h["front_face"][hit] == 0)


def test_furnace_closed_sphere():
    """Inside a closed emitting-free Lambertian sphere lit by a constant background seen through nothing: a path that
    starts inside a sphere of albedo rho bounces forever; with emission Le on an inner light sphere the radiance is
    finite.  Simpler closed form used here: camera inside a sphere whose material is DiffuseLight(Le) -> every pixel = Le."""
    light = rt.DiffuseLight(rt.ConstantTexture((0.25, 0.5, 1.0)))
    sb = rt.World(rt.HitableList([rt.Sphere((0, 0, 0), 10.0, light)]), rt.Camera(((0, 0, 0), (0, 0, -1)), (0, 1, 0), 90, 1.0, 0.0, 1.0)).builder
    p = rt.make_params(16, 16, 4)
    acc, ext, shd, _ = OracleScene(sb).render(sb.camera, p, n_jobs=4, n_threads=2)
    np.testing.assert_allclose(acc / 4, np.broadcast_to([0.25, 0.5, 1.0], (16, 16, 3)), rtol=1e-12)
    assert ext == 16 * 16 * 4 and shd == 0


def test_white_furnace_lambertian_albedo_series():
    """Camera inside a Lambertian sphere (albedo rho) with a constant white background that is never reached, plus the
    LIGHT_SCATTERS-free loop: radiance = 0 (all paths die at depth 50); with max_depth = 0 every path stops at once."""
    lam = rt.Lambertian(rt.ConstantTexture((0.5, 0.5, 0.5)))
    sb = rt.World(rt.HitableList([rt.Sphere((0, 0, 0), 10.0, lam)]), rt.Camera(((0, 0, 0), (0, 0, -1)), (0, 1, 0), 90, 1.0, 0.0, 1.0)).builder
    p = rt.make_params(8, 8, 2, max_depth=5)
    acc, ext, _, _ = OracleScene(sb).render(sb.camera, p, n_jobs=2, n_threads=2)
    assert np.all(acc == 0) and ext == 8 * 8 * 2 * 6  # 5 scattered hits + the hit that finds the budget empty (photon_mapper.rs:335-339)


def test_three_estimators_agree_in_the_mean():
    """PATH, MIXTURE and NEE are unbiased estimators of the same integral (SURVEY.md §8c cross-check)."""
    sb = scene_builder("cornell_book3")
    orc = OracleScene(sb)
    means = []
    for integ in (F.INTEGRATOR_PATH, F.INTEGRATOR_MIXTURE, F.INTEGRATOR_NEE):
        p = rt.copy_params(sb.params, width=32, height=32, spp_count=96 if integ else 256, integrator=integ)
        acc, _, _, _ = orc.render(sb.camera, p)
        means.append(acc.mean(axis=(0, 1)) / p.spp_count)
    means = np.array(means)
    assert np.all(np.abs(means[1] - means[2]) / means[2] < 0.05), means
    assert np.all(np.abs(means[0] - means[2]) / means[2] < 0.12), means
