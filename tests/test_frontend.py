"""Host logic: the mirror of the reference's construction API flattens into the tables of include/b200rt.h."""
import ctypes as C
import json
import os

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from scenes_util import DATA, scene_builder


def test_reference_cornell_scene_tables():
    """scene.rs:16-112: 5 walls + light + 2 spheres + Transform(Mesh(cube.obj)) (12 triangles) + Cube."""
    d = scene_builder("cornell_ref").desc()
    types = [d.prims[i].type for i in range(d.n_prims)]
    assert d.n_prims == 5 + 1 + 2 + 12 + 1
    assert types.count(F.PRIM_TRIANGLE) == 12 and types.count(F.PRIM_BOX) == 1 and types.count(F.PRIM_SPHERE) == 2
    assert d.n_instances == 1 and d.n_lights == 1
    tri = [d.prims[i] for i in range(d.n_prims) if d.prims[i].type == F.PRIM_TRIANGLE]
    assert all(t.instance == 0 for t in tri)
    m = np.array(d.instances[0].m[:]).reshape(3, 4)  # M = T(100,50,100) * S(50)
    np.testing.assert_allclose(m, [[50, 0, 0, 100], [0, 50, 0, 50], [0, 0, 50, 100]])
    assert d.prims[d.light_prims[0]].type == F.PRIM_XZ_RECT
    # one table row per Arc<dyn Material>: `white` is shared by three walls, the mesh and the cube
    assert d.n_materials == 6


def test_cube_obj_loader_matches_tobj_semantics():
    """mesh.rs:148-199 on data/mesh/cube.obj: 12 triangles, 6 distinct (flat) normals, f32 positions."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    w = rt.World(rt.HitableList([rt.Mesh.load_obj(os.path.join(DATA, "mesh", "cube.obj"), white)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    d = w.builder.desc()
    assert d.n_prims == 12 and d.n_tri_normals == 12
    normals = {tuple(np.round(d.tri_normals[9 * i:9 * i + 3], 6)) for i in range(12)}
    assert len(normals) == 6
    pos = np.array([d.prims[i].p[:9] for i in range(12)])
    assert np.abs(pos).max() <= 1.000002


def test_obj_with_several_models_and_texcoords_loads_like_mesh_rs(tmp_path):
    """Mesh::load_obj takes `models[0]` only (mesh.rs:159) and Triangle::hit hands HitRecord::new the uv (0.0, 0.0) — the texcoord
    interpolation is a commented-out todo!() upstream (mesh.rs:66-68, 103-128).  A two-object OBJ with vt lines and v/vt/vn faces:
    the first object's triangles only, texture coordinates read past."""
    obj = tmp_path / "two.obj"
    obj.write_text("""o first
v 0 0 0
v 1 0 0
v 0 1 0
v 1 1 0
vt 0 0
vt 1 0
vt 0 1
vt 1 1
vn 0 0 1
f 1/1/1 2/2/1 3/3/1
f 2/2/1 4/4/1 3/3/1
o second
v 0 0 5
v 1 0 5
v 0 1 5
f 5//1 6//1 7//1
""")
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    w = rt.World(rt.HitableList([rt.Mesh.load_obj(str(obj), white)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    d = w.builder.desc()
    assert d.n_prims == 2
    z = np.array([d.prims[i].p[:9] for i in range(2)]).reshape(2, 3, 3)[..., 2]
    assert np.all(z == 0)  # nothing of the second object


def test_obj_without_normals_is_an_error_like_the_reference_panic():
    """bun315.obj has no vn lines: Triangle::hit would index an empty normals Vec (mesh.rs:62-64)."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    with pytest.raises(rt.B200RTError) as e:
        rt.World(rt.HitableList([rt.Mesh.load_obj(os.path.join(DATA, "mesh", "bun315.obj"), white)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    assert "normals" in str(e.value)
    w = rt.World(rt.HitableList([rt.Mesh.load_obj(os.path.join(DATA, "mesh", "bun315.obj"), white, make_normals=True)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    assert w.builder.desc().n_prims == 4968
    w = rt.World(rt.HitableList([rt.Mesh.load_obj(os.path.join(DATA, "mesh", "bun315.obj"), white, make_normals=True, subdivide=1)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    assert w.builder.desc().n_prims == 4 * 4968


def test_missing_obj_and_bad_ids():
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    with pytest.raises(rt.B200RTError) as e:
        rt.World(rt.HitableList([rt.Mesh.load_obj("/nonexistent.obj", white)]), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1))
    assert e.value.status == F.ERR_IO
    sb = rt.SceneBuilder()
    assert F.lib().b200rt_builder_sphere(sb.handle, F.d3((0, 0, 0)), 1.0, 7) < 0  # unknown material id
    with pytest.raises(rt.B200RTError):
        rt.Camera(((0, 0, 0), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1)  # look_from == look_at: unit() of a zero vector (vec3.rs:88)


@pytest.mark.parametrize("name,n,flat", [("scene_10", 25, 0), ("scene_200_no_bvh", 405, 1), ("scene_500", 1005, 0)])
def test_json_scenes(name, n, flat):
    sb = rt.SceneBuilder().load_json(open(os.path.join(DATA, name + ".json")).read())
    d = sb.desc()
    assert d.n_prims == n and (d.flags & F.SCENE_FLAT_LIST) == flat
    # the five fixed spheres of Appendix B
    assert tuple(np.round(d.prims[0].p[:4], 4)) == (0.0, -100.5, -1.0, 100.0)
    mats = [d.materials[d.prims[i].material].type for i in range(5)]
    assert mats == [F.MAT_LAMBERTIAN, F.MAT_DIELECTRIC, F.MAT_DIFFUSE_LIGHT, F.MAT_METAL, F.MAT_DIELECTRIC]
    assert d.textures[d.materials[d.prims[0].material].texture].type == F.TEX_CHECKER
    cam = sb.camera
    np.testing.assert_allclose(cam.origin[:], [-6, 2, -6])
    assert abs(cam.lens_radius - 0.05) < 1e-7


@pytest.mark.parametrize("name", ["scene_10", "scene_200_no_bvh"])
def test_yaml_twin_builds_the_same_tables(name):
    """README.md:86-89: the .yaml files carry the same objects as the .json files -> identical flat tables and camera."""
    a = rt.SceneBuilder().load_json(open(os.path.join(DATA, name + ".json")).read())
    b = rt.SceneBuilder().load_yaml(open(os.path.join(DATA, name + ".yaml")).read())
    da, db = a.desc(), b.desc()
    assert (da.n_prims, da.n_materials, da.n_textures, da.flags) == (db.n_prims, db.n_materials, db.n_textures, db.flags)
    for i in range(da.n_prims):
        pa, pb = da.prims[i], db.prims[i]
        assert (pa.type, pa.material, pa.instance) == (pb.type, pb.material, pb.instance)
        assert list(pa.p) == list(pb.p)
    for i in range(da.n_materials):
        assert (da.materials[i].type, da.materials[i].texture, da.materials[i].param) == (db.materials[i].type, db.materials[i].texture, db.materials[i].param)
    for i in range(da.n_textures):
        assert (da.textures[i].type, list(da.textures[i].c0), list(da.textures[i].c1)) == (db.textures[i].type, list(db.textures[i].c0), list(db.textures[i].c1))
    for f in ("origin", "lower_left_corner", "horizontal", "vertical"):
        assert list(getattr(a.camera, f)) == list(getattr(b.camera, f))
    assert a.camera.lens_radius == b.camera.lens_radius


def test_yaml_errors_and_plain_style():
    with pytest.raises(rt.B200RTError) as e:
        rt.SceneBuilder().load_yaml("objects:\n  type: Torus\n")
    assert e.value.status == F.ERR_PARSE
    with pytest.raises(rt.B200RTError):
        rt.SceneBuilder().load_yaml("objects\n")
    # a sequence kept at its key's indentation (plain YAML style) and a comment line
    text = "\n".join([
        "# one sphere", "objects:", "  type: HitableList", "  items:", "  - type: Sphere", "    center: {}".replace("{}", ""),
        "      x: 0", "      y: 0", "      z: -1", "    radius: 0.5", "    material:", "      type: Dielectric", "      ref_idx: 1.5",
        "camera:", "  look_from:", "    x: 0", "    y: 0", "    z: 3", "  look_at:", "    x: 0", "    y: 0", "    z: 0", "  vup:", "    x: 0", "    y: 1", "    z: 0",
        "  vfov: 40", "  aspect: 1", "  aperture: 0", "  focus_dist: 3", ""])
    sb = rt.SceneBuilder().load_yaml(text)
    d = sb.desc()
    assert d.n_prims == 1 and tuple(d.prims[0].p[:4]) == (0.0, 0.0, -1.0, 0.5)
    assert d.materials[d.prims[0].material].type == F.MAT_DIELECTRIC


def test_json_errors():
    with pytest.raises(rt.B200RTError) as e:
        rt.SceneBuilder().load_json('{"objects": {"type": "Torus"}}')
    assert e.value.status == F.ERR_PARSE
    with pytest.raises(rt.B200RTError):
        rt.SceneBuilder().load_json('{"objects": ')


def test_transform_composition_and_rotated_shapes():
    """M = T*S*Rx*Ry*Rz (transform.rs:28-106).  Every shape keeps its own type under any Transform (transform.rs:17-165 takes
    any rotation, scale: Vec3 and any child): a rotated Cube stays ONE box primitive with an instance matrix (rectangle uv and
    the six-rect tie rule of cube.rs:63-66 survive), and a sphere under a non-uniform scale is accepted (object-space test)."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    items = rt.HitableList([
        rt.Transform((10, 20, 30), (2, 2, 2), (1, 2, 3), rt.Sphere((0, 0, 0), 1, white)),
        rt.Transform((0, 15, 0), (1, 1, 1), (265, 0, 295), rt.Cube((0, 0, 0), (165, 330, 165), white)),
        rt.Transform((0, 0, 0), (2, 3, 4), (0, 0, 0), rt.Cube((0, 0, 0), (1, 1, 1), white)),
        rt.Transform((30, 0, 0), (1, 1, 1), (0, 0, 0), rt.XZRectangle((0, 0), (1, 1), 0.5, white)),
        rt.Transform((0, 0, 0), (1, 2, 1), (0, 0, 0), rt.Sphere((0, 0, 0), 1, white)),  # an ellipsoid in world space
        rt.Transform((0, 0, 45), (1, 1, 1), (5, 0, 0), rt.ConstantMedium(0.5, rt.Cube((0, 0, 0), (1, 1, 1), white), rt.Isotropic(rt.ConstantTexture((1, 1, 1))))),
        rt.Sphere((0, 0, 0), -0.9, white),  # negative radius: the hollow-sphere idiom (sphere.rs:44 divides by r)
    ])
    d = rt.World(items, rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1)).builder.desc()
    types = [d.prims[i].type for i in range(d.n_prims)]
    assert types == [F.PRIM_SPHERE, F.PRIM_BOX, F.PRIM_BOX, F.PRIM_XZ_RECT, F.PRIM_SPHERE, F.PRIM_MEDIUM_BOX, F.PRIM_SPHERE]
    assert [d.prims[i].instance for i in range(d.n_prims)] == [0, 1, 2, 3, 4, 5, -1]
    rx, ry, rz = np.radians([10, 20, 30])
    Rx = np.array([[1, 0, 0], [0, np.cos(rx), -np.sin(rx)], [0, np.sin(rx), np.cos(rx)]])
    Ry = np.array([[np.cos(ry), 0, np.sin(ry)], [0, 1, 0], [-np.sin(ry), 0, np.cos(ry)]])
    Rz = np.array([[np.cos(rz), -np.sin(rz), 0], [np.sin(rz), np.cos(rz), 0], [0, 0, 1]])
    want = np.hstack([2 * Rx @ Ry @ Rz, [[1], [2], [3]]])
    np.testing.assert_allclose(np.array(d.instances[0].m[:]).reshape(3, 4), want, atol=1e-15)
    st = rt.compile_stats(d)  # World::new's host half accepts all of it (no CUDA needed)
    assert st["bvh_nodes"] >= 1 and st["stack_needed"] <= st["stack_capacity"]
    sb = rt.SceneBuilder()
    m = sb._id(F.lib().b200rt_builder_material(sb.handle, F.MAT_LAMBERTIAN, sb._id(F.lib().b200rt_builder_texture_constant(sb.handle, 1, 1, 1)), 0))
    sb._id(F.lib().b200rt_builder_sphere(sb.handle, F.d3((0, 0, 0)), 0.0, m))
    with pytest.raises(rt.B200RTError) as e:  # radius 0 divides by zero upstream
        rt.compile_stats(sb)
    assert e.value.status == F.ERR_INVALID


def test_bvh_depth_is_bounded_by_the_traversal_stack():
    """The device keeps a per-thread traversal stack and pushes up to 3 entries per BVH4 level.  A chain of spheres of
    geometrically growing size along a line makes a plain SAH build peel one primitive per level (binary depth ~n): World::new
    must come back with a tree whose deepest path fits the stack (rebuilt under a depth budget), never with a silent overflow."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    for n, ratio in ((200, 1.6), (400, 1.2), (3000, 1.01)):
        items, x, r = [], 0.0, 1e-3
        for _ in range(n):
            items.append(rt.Sphere((x + r, 0, 0), r, white))
            x += 2 * r
            r *= ratio
            if r > 1e30:
                break
        st = rt.compile_stats(rt.World(rt.HitableList(items), rt.Camera(((0, 0, 5), (0, 0, 0)), (0, 1, 0), 40, 1, 0, 1)).builder)
        assert st["stack_needed"] == 3 * st["bvh4_depth"] and st["stack_needed"] <= st["stack_capacity"], (n, st)
    st = rt.compile_stats(scene_builder("book2"))
    assert st["bvh4_depth"] <= 10  # the budget never touches an ordinary scene


def test_bvh_quality_size_axis_and_dense_box(monkeypatch):
    """Host only.  A mesh in a room: the builder's size axis puts the walls on one side of the root and the mesh on the other, which
    lowers the surface-area estimate of the tree and exposes the mesh's box as the scene's dense box; the book scenes have none."""
    sb = scene_builder("mesh_small")
    q = rt.bvh_quality(sb)
    assert q["dense_box"] is not None
    lo, hi = (np.array(v) for v in q["dense_box"])
    assert np.all(lo > -1.0) and np.all(hi < 556.0) and np.all(hi > lo)
    ext = hi - lo
    assert 2 * (ext[0] * ext[1] + ext[1] * ext[2] + ext[2] * ext[0]) <= 0.5 * 6 * 555.1 ** 2
    monkeypatch.setenv("B200RT_BVH_SIZE_SPLIT", "0")
    q0 = rt.bvh_quality(sb)
    monkeypatch.delenv("B200RT_BVH_SIZE_SPLIT")
    assert q0["dense_box"] is None and q["node_steps"] < 0.9 * q0["node_steps"] and q["prim_tests"] <= 1.05 * q0["prim_tests"]
    for name in ("book1", "book2", "cornell_book3", "json:scene_500"):
        assert rt.bvh_quality(scene_builder(name))["dense_box"] is None, name
    with pytest.raises(rt.B200RTError):
        rt.bvh_quality(rt.SceneBuilder())  # BVHNode::new on an empty list


def test_dense_box_of_a_cluster_in_a_room():
    """Host only, through the construction API: five walls around 6,000 small spheres in one corner.  The dense box is the
    cluster's; a room that the cluster fills has none; neither has a room without a cluster."""
    white = rt.Lambertian(rt.ConstantTexture((1, 1, 1)))
    cam = rt.Camera(((50, 50, -150), (50, 50, 50)), (0, 1, 0), 40, 1, 0, 1)
    walls = [rt.YZRectangle((0, 0), (100, 100), 0, white), rt.YZRectangle((0, 0), (100, 100), 100, white), rt.XZRectangle((0, 0), (100, 100), 0, white),
             rt.XZRectangle((0, 0), (100, 100), 100, white), rt.XYRectangle((0, 0), (100, 100), 100, white)]
    rng = np.random.default_rng(5)

    def cluster(lo, hi, n=6000):
        c = rng.uniform(lo, hi, size=(n, 3))
        return [rt.Sphere(tuple(map(float, x)), 0.2, white) for x in c]

    q = rt.bvh_quality(rt.World(rt.HitableList(walls + cluster(10, 40)), cam).builder)
    assert q["dense_box"] is not None
    lo, hi = (np.array(v) for v in q["dense_box"])
    assert np.all(lo > 9.5) and np.all(lo < 10.5) and np.all(hi > 39.5) and np.all(hi < 40.5)
    assert rt.bvh_quality(rt.World(rt.HitableList(walls + cluster(1, 99)), cam).builder)["dense_box"] is None  # 94 % of the room's surface
    assert rt.bvh_quality(rt.World(rt.HitableList(walls), cam).builder)["dense_box"] is None


@pytest.mark.parametrize("name", ["mesh", "book2"])
def test_bvh_does_not_depend_on_the_builder_threads(name, monkeypatch):
    """Host only.  The builder assigns node slots up front and builds the halves of a large node on two threads: same tree
    (node count, depth, surface-area sums, dense box) for 1, 3 and 16 threads."""
    sb = scene_builder(name)
    seen = []
    for threads in ("1", "3", "16"):
        monkeypatch.setenv("B200RT_BVH_THREADS", threads)
        seen.append((rt.compile_stats(sb), rt.bvh_quality(sb)))
    assert seen[0] == seen[1] == seen[2]


def test_scene_500_yaml_builds_the_same_tables_as_its_json_twin():
    """README.md:86-89: data/scene_500.yaml (894 KB, 1,005 spheres, 999 BVHNodes) carries the same objects as scene_500.json."""
    a = rt.SceneBuilder().load_json(open(os.path.join(DATA, "scene_500.json")).read())
    b = rt.SceneBuilder().load_yaml(open(os.path.join(DATA, "scene_500.yaml")).read())
    da, db = a.desc(), b.desc()
    assert da.n_prims == db.n_prims == 1005 and da.n_materials == db.n_materials and da.n_textures == db.n_textures and da.flags == db.flags
    assert bytes(C.string_at(da.prims, 64 * da.n_prims)) == bytes(C.string_at(db.prims, 64 * db.n_prims))
    assert bytes(C.string_at(da.materials, 16 * da.n_materials)) == bytes(C.string_at(db.materials, 16 * db.n_materials))
    assert bytes(C.string_at(da.textures, 48 * da.n_textures)) == bytes(C.string_at(db.textures, 48 * db.n_textures))
    assert bytes(a.camera) == bytes(b.camera)


def test_camera_new_matches_reference_formula():
    """camera.rs:24-55 evaluated in numpy."""
    c = rt.Camera(((13, 2, 3), (0, 0, 0)), (0, 1, 0), 20, 16 / 9, 0.1, 10.0).c
    w = np.array([13, 2, 3.0]); w /= np.linalg.norm(w)
    u = np.cross([0, 1, 0], w); u /= np.linalg.norm(u)
    v = np.cross(w, u)
    h = np.tan(np.radians(20) / 2)
    hor, ver = 10 * (16 / 9) * 2 * h * u, 10 * 2 * h * v
    np.testing.assert_allclose(c.horizontal[:], hor, rtol=1e-6)
    np.testing.assert_allclose(c.vertical[:], ver, rtol=1e-6)
    np.testing.assert_allclose(c.lower_left_corner[:], np.array([13, 2, 3]) - hor / 2 - ver / 2 - 10 * w, rtol=1e-6)
    assert abs(c.lens_radius - 0.05) < 1e-8


def test_canonical_scenes_are_deterministic_in_the_seed():
    a, b, c = scene_builder("book1", 1).desc(), scene_builder("book1", 1).desc(), scene_builder("book1", 2).desc()
    pa = np.array([a.prims[i].p[:4] for i in range(a.n_prims)])
    pb = np.array([b.prims[i].p[:4] for i in range(b.n_prims)])
    pc = np.array([c.prims[i].p[:4] for i in range(min(a.n_prims, c.n_prims))])
    assert np.array_equal(pa, pb) and not np.array_equal(pa[:len(pc)], pc)
    d = scene_builder("book2").desc()
    types = [d.prims[i].type for i in range(d.n_prims)]
    assert types.count(F.PRIM_BOX) == 400 and types.count(F.PRIM_MEDIUM_SPHERE) == 2 and types.count(F.PRIM_MOVING_SPHERE) == 1
    assert sum(1 for i in range(d.n_prims) if d.prims[i].instance >= 0) == 1000 and d.perlin and d.n_images == 1
