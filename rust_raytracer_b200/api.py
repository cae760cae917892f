"""Python mirror of the reference's scene-construction API over the C ABI (tests and bench use this).

Names, constructor arguments and nesting follow raytracer/src/{objects/*,material,camera,world,scene}.rs; each
type gains one method, `flatten(builder)`, exactly like the C++ mirror (csrc/frontend.hpp) and the Rust
binding sketched in INTEGRATION.md.  All geometry work happens in libb200rt.so.
"""
import ctypes as C

import numpy as np

from . import _ffi as F


class B200RTError(RuntimeError):
    def __init__(self, status, message):
        super().__init__("b200rt status %d: %s" % (status, message))
        self.status = status


class Vec3(tuple):
    def __new__(cls, x=0.0, y=0.0, z=0.0):
        return super().__new__(cls, (float(x), float(y), float(z)))

    @staticmethod
    def ones():
        return Vec3(1, 1, 1)

    @staticmethod
    def zero():
        return Vec3(0, 0, 0)

    def __mul__(self, s):
        return Vec3(self[0] * s, self[1] * s, self[2] * s)


class SceneBuilder:
    """Owns a b200rt_builder; `desc()` returns the flat tables (valid while the builder lives)."""

    def __init__(self):
        self._h = C.c_void_p()
        rc = F.lib().b200rt_builder_create(C.byref(self._h))
        if rc:
            raise B200RTError(rc, "builder_create")
        self._cache = {}
        self.camera = F.Camera()
        self.params = F.Params()

    def __del__(self):
        try:
            if self._h:
                F.lib().b200rt_builder_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def _id(self, v):
        if v < 0:
            raise B200RTError(-v, F.lib().b200rt_builder_error(self._h).decode())
        return v

    def _ok(self, rc):
        if rc:
            raise B200RTError(rc, F.lib().b200rt_builder_error(self._h).decode())

    @property
    def handle(self):
        return self._h

    def desc(self):
        d = F.SceneDesc()
        self._ok(F.lib().b200rt_builder_finish(self._h, C.byref(d)))
        d._builder = self  # the tables live in the builder: keep it alive as long as the desc
        return d

    def set_flags(self, flags):
        self._ok(F.lib().b200rt_builder_set_flags(self._h, flags))

    def load_json(self, text):
        data = text.encode() if isinstance(text, str) else bytes(text)
        self._ok(F.lib().b200rt_builder_load_json(self._h, data, len(data), C.byref(self.camera)))
        return self

    def load_yaml(self, text):
        data = text.encode() if isinstance(text, str) else bytes(text)
        self._ok(F.lib().b200rt_builder_load_yaml(self._h, data, len(data), C.byref(self.camera)))
        return self

    def make_scene(self, name, seed=1, aux_path=None, earth=None, aspect=0.0):
        ptr, w, h = None, 0, 0
        if earth is not None:
            earth = np.ascontiguousarray(earth, dtype=np.uint8)
            h, w = earth.shape[0], earth.shape[1]
            ptr = earth.ctypes.data_as(C.c_void_p)
        self._ok(F.lib().b200rt_builder_make_scene(self._h, name.encode(), seed, aux_path.encode() if aux_path else None, ptr, w, h, aspect,
                                                   C.byref(self.camera), C.byref(self.params)))
        return self


# ---- textures (material.rs:48-84) ------------------------------------------------------------------------
class ConstantTexture:
    def __init__(self, color):
        self.color = Vec3(*color)

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_texture_constant(sb.handle, *self.color))


class CheckerTexture:
    def __init__(self, t0, t1):
        self.t0, self.t1 = t0, t1

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_texture_checker(sb.handle, F.d3(self.t0.color), F.d3(self.t1.color)))


class ImageTexture:
    def __init__(self, rgb8):
        self.rgb = np.ascontiguousarray(rgb8, dtype=np.uint8)

    def flatten(self, sb):
        h, w = self.rgb.shape[:2]
        return sb._id(F.lib().b200rt_builder_texture_image(sb.handle, self.rgb.ctypes.data_as(C.c_void_p), w, h))


class NoiseTexture:
    def __init__(self, scale):
        self.scale = scale

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_texture_noise(sb.handle, self.scale))


# ---- materials (material.rs:86-231) ----------------------------------------------------------------------
class _Material:
    TYPE = -1

    def __init__(self, texture, param=0.0):
        self.texture, self.param = texture, param

    def flatten(self, sb):
        key = id(self)
        if key not in sb._cache:
            sb._cache[key] = sb._id(F.lib().b200rt_builder_material(sb.handle, self.TYPE, self.texture.flatten(sb), self.param))
        return sb._cache[key]


class Lambertian(_Material):
    TYPE = F.MAT_LAMBERTIAN

    def __init__(self, albedo):
        super().__init__(albedo)


class Metal(_Material):
    TYPE = F.MAT_METAL

    def __init__(self, albedo, fuzz):
        super().__init__(albedo, fuzz)


class Dielectric(_Material):
    TYPE = F.MAT_DIELECTRIC

    def __init__(self, ir, albedo):
        super().__init__(albedo, ir)


class DiffuseLight(_Material):
    TYPE = F.MAT_DIFFUSE_LIGHT

    def __init__(self, emit):
        super().__init__(emit)


class Isotropic(_Material):
    TYPE = F.MAT_ISOTROPIC

    def __init__(self, albedo):
        super().__init__(albedo)


# ---- hitables (objects/*.rs) -----------------------------------------------------------------------------
class Sphere:
    def __init__(self, center, radius, material):
        self.center, self.radius, self.material = Vec3(*center), radius, material

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_sphere(sb.handle, F.d3(self.center), self.radius, self.material.flatten(sb)))


class MovingSphere:
    def __init__(self, center0, center1, time0, time1, radius, material):
        self.c0, self.c1, self.t0, self.t1, self.radius, self.material = Vec3(*center0), Vec3(*center1), time0, time1, radius, material

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_moving_sphere(sb.handle, F.d3(self.c0), F.d3(self.c1), self.t0, self.t1, self.radius, self.material.flatten(sb)))


class _AxisRectangle:
    AXIS = -1

    def __init__(self, p0, p1, k, material):
        self.p0, self.p1, self.k, self.material = p0, p1, k, material

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_rect(sb.handle, self.AXIS, self.p0[0], self.p0[1], self.p1[0], self.p1[1], self.k, self.material.flatten(sb)))


class XYRectangle(_AxisRectangle):  # {xy0, xy1, z, material}
    AXIS = 2


class XZRectangle(_AxisRectangle):  # {xz0, xz1, y, material}
    AXIS = 1


class YZRectangle(_AxisRectangle):  # {yz0, yz1, x, material}
    AXIS = 0


class Cube:
    def __init__(self, box_min, box_max, material):
        self.box_min, self.box_max, self.material = Vec3(*box_min), Vec3(*box_max), material

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_cube(sb.handle, F.d3(self.box_min), F.d3(self.box_max), self.material.flatten(sb)))


class Triangle:
    def __init__(self, pa, pb, pc, na, nb, nc, material):
        self.v = (pa, pb, pc, na, nb, nc)
        self.material = material

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_triangle(sb.handle, *[F.d3(x) for x in self.v], self.material.flatten(sb)))


class Mesh:
    def __init__(self, obj_file, material, make_normals=False, subdivide=0):
        self.obj_file, self.material, self.make_normals, self.subdivide = obj_file, material, make_normals, subdivide

    @staticmethod
    def load_obj(obj_file, material, make_normals=False, subdivide=0):
        return Mesh(obj_file, material, make_normals, subdivide)

    def flatten(self, sb):
        return sb._id(F.lib().b200rt_builder_mesh_obj(sb.handle, self.obj_file.encode(), self.material.flatten(sb), int(self.make_normals), self.subdivide))


class HitableList(list):
    def flatten(self, sb):
        first = -1
        for h in self:
            i = h.flatten(sb)
            first = i if first < 0 else first
        return first


class BVHNode(HitableList):
    """BVHNode::new(src_objects) (bvh.rs:60-83): forwards its children; the device BVH is rebuilt wide."""

    @staticmethod
    def new(objects):
        return BVHNode(objects)


class Transform:
    def __init__(self, rotate_in_degree, scale, translate, obj):
        self.r, self.s, self.t, self.obj = rotate_in_degree, scale, translate, obj

    def flatten(self, sb):
        sb._ok(F.lib().b200rt_builder_push_transform(sb.handle, F.d3(self.r), F.d3(self.s), F.d3(self.t)))
        try:
            return self.obj.flatten(sb)
        finally:
            F.lib().b200rt_builder_pop_transform(sb.handle)


class ConstantMedium:
    def __init__(self, density, boundary, phase_function):
        self.density, self.boundary, self.phase = density, boundary, phase_function

    def flatten(self, sb):
        b = self.boundary
        if isinstance(b, Sphere):
            return sb._id(F.lib().b200rt_builder_medium_sphere(sb.handle, F.d3(b.center), b.radius, self.density, self.phase.flatten(sb)))
        if isinstance(b, Cube):
            return sb._id(F.lib().b200rt_builder_medium_cube(sb.handle, F.d3(b.box_min), F.d3(b.box_max), self.density, self.phase.flatten(sb)))
        raise B200RTError(F.ERR_UNSUPPORTED, "ConstantMedium boundary must be a Sphere or a Cube")


class Camera:
    """Camera::new((look_from, look_at), vup, vfov, aspect_ratio, aperture, focus_dist) (camera.rs:24-31)."""

    def __init__(self, look_from_to, vup, vfov, aspect_ratio, aperture, focus_dist, time0=0.0, time1=0.0):
        self.c = F.Camera()
        rc = F.lib().b200rt_camera_make(F.d3(look_from_to[0]), F.d3(look_from_to[1]), F.d3(vup), vfov, aspect_ratio, aperture, focus_dist, time0, time1, C.byref(self.c))
        if rc:
            raise B200RTError(rc, "Camera::new: degenerate look_from / look_at / vup")


class World:
    """World::new(hitable_list, cam, lights) (world.rs:15-25)."""

    def __init__(self, hitable_list, cam, lights=()):
        self.builder = SceneBuilder()
        ids = {}
        for h in hitable_list:
            ids[id(h)] = h.flatten(self.builder)
        for l in lights:  # a hitable, or (hitable, scale): the `scale` of XZRectLight::new / SphereDiffuseLight::new (light.rs:71,131)
            l, scale = l if isinstance(l, tuple) else (l, 1.0)
            if id(l) not in ids:
                raise B200RTError(F.ERR_INVALID, "World::new: a light must also be in the hitable list (scene.rs:86,110)")
            self.builder._ok(F.lib().b200rt_builder_mark_light_scaled(self.builder.handle, ids[id(l)], float(scale)))
        self.cam = cam
        self.builder.camera = cam.c
