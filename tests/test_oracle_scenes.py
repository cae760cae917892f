"""The oracle builds the canonical scenes and the JSON files ITSELF (oracle/oracle_scenes.inc: the reference's object graph —
Transform over Cube over six rectangles, Mesh = BVHNode over Triangles, nested BVHNode::new, World::new), never from the
product's tables.  Here, on the CPU, the two constructions are set against each other inside the oracle: the graph the
oracle made on its own and the graph rebuilt from what libb200rt.so's builder / loaders / scene generators flattened must
answer every ray alike.  A flattening bug (transform composition, light marking, JSON walk order, a generator) shows up
here instead of cancelling out between oracle and device."""
import ctypes as C
import os

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
import oracle_ffi
from oracle_ffi import OracleScene
from scenes_util import scene_builder, oracle_scene, ray_batch

NAMES = ["book1", "cornell_ref", "cornell_book3", "book2", "mesh_small", "json:scene_10", "json:scene_200_no_bvh", "json:scene_500"]


def test_abi_struct_mirrors_have_the_product_layouts():
    for a, b in ((oracle_ffi.oracle_py.Camera, F.Camera), (oracle_ffi.oracle_py.Params, F.Params)):
        assert C.sizeof(a) == C.sizeof(b)
        assert [(n, C.sizeof(t)) for n, t in a._fields_] == [(n, C.sizeof(t)) for n, t in b._fields_]


def test_oracle_binding_does_not_import_the_product():
    import re
    src = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle", "oracle_py.py")).read()
    code = src.split('"""', 2)[2]  # past the module docstring
    assert not re.search(r"^\s*(import|from)\s+rust_raytracer_b200", code, re.M) and "libb200rt" not in code


@pytest.mark.parametrize("name", NAMES)
def test_own_graph_equals_flattened_graph(built, name):
    sb = scene_builder(name)
    own, flat = oracle_scene(name), OracleScene(sb)
    assert own.n_prims == sb.desc().n_prims  # a leaf object's id = its position in construction order, on both sides
    assert bytes(own.camera) == bytes(sb.camera)  # Camera::new (camera.rs:24-55), computed in f64 then rounded, twice
    if not name.startswith("json:"):
        assert bytes(own.params) == bytes(sb.params)
    rng = np.random.default_rng(11)
    o, d, seeds = ray_batch(sb, 3000 if name == "mesh_small" else 40000, rng, cam_res=(40, 30) if name == "mesh_small" else (96, 64))
    a, b = own.hit_batch(o, d, seeds), flat.hit_batch(o, d, seeds)
    assert np.array_equal(a["prim"] >= 0, b["prim"] >= 0)
    hit = a["prim"] >= 0
    assert hit.sum() > 1000
    assert np.array_equal(a["t"][hit], b["t"][hit])  # the same f64 arithmetic on the same f32-rounded inputs: bit-equal t
    tie = hit & (a["prim"] != b["prim"])  # two surfaces at exactly the same t: bvh.rs:91-101 resolves it by (random) topology
    assert tie.sum() <= 2e-3 * hit.sum(), (name, int(tie.sum()))
    same = hit & ~tie
    for f in ("nx", "ny", "nz", "u", "v", "front_face", "px", "py", "pz"):
        assert np.array_equal(a[f][same], b[f][same]), (name, f)


def test_own_book2_images_agree(built):
    """capture_image over both graphs with the same seeds: identical paths, hence the same image up to ties."""
    sb = scene_builder("book2")
    own, flat = oracle_scene("book2"), OracleScene(sb)
    p = rt.copy_params(sb.params, width=24, height=24, spp_count=4)
    a, ea, _, _ = own.render(sb.camera, p, n_jobs=4, n_threads=4)
    b, eb, _, _ = flat.render(sb.camera, p, n_jobs=4, n_threads=4)
    assert ea == eb or abs(ea - eb) < 0.01 * ea
    assert abs(a.mean() - b.mean()) <= 0.02 * b.mean() + 1e-9


def test_tight_triangle_boxes_change_no_answer(built):
    """mesh.rs:33-42 pads every triangle box by 0.1; the padding only affects how much the BVH prunes."""
    sb = scene_builder("mesh_small")
    rng = np.random.default_rng(5)
    o, d, seeds = ray_batch(sb, 1500, rng, cam_res=(32, 24))
    a = oracle_scene("mesh_small").hit_batch(o, d, seeds)
    b = oracle_scene("mesh_small", tri_pad=1e-6).hit_batch(o, d, seeds)
    assert np.array_equal(a["prim"] >= 0, b["prim"] >= 0)
    hit = a["prim"] >= 0
    assert np.array_equal(a["t"][hit], b["t"][hit])
    assert (a["prim"][hit] != b["prim"][hit]).mean() < 2e-3
