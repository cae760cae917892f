"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE (only tests/, smoke() and bench.py's CPU legs load it).

Standalone on purpose: it declares its own mirrors of the public ABI structs (include/b200rt.h) and imports nothing
from rust_raytracer_b200, so `bench.py --impl reference` maps liboracle.so and nothing else.  tests/test_oracle_scenes.py
checks that these mirrors have the layouts of the product's own ctypes classes.
"""
import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
ORACLE_PATH = os.environ.get("ORACLE_LIB") or os.path.join(HERE, "liboracle.so")  # ORACLE_LIB: an instrumented build (make -C oracle asan)


class Camera(C.Structure):  # b200rt_camera
    _fields_ = [("origin", C.c_float * 3), ("lower_left_corner", C.c_float * 3), ("horizontal", C.c_float * 3), ("vertical", C.c_float * 3),
                ("u", C.c_float * 3), ("v", C.c_float * 3), ("w", C.c_float * 3), ("lens_radius", C.c_float), ("time0", C.c_float), ("time1", C.c_float)]


class Params(C.Structure):  # b200rt_params
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("spp_begin", C.c_int32), ("spp_count", C.c_int32), ("max_depth", C.c_int32),
                ("integrator", C.c_int32), ("pipeline", C.c_int32), ("background_mode", C.c_int32), ("t_min", C.c_float),
                ("background", C.c_float * 3), ("flags", C.c_uint32), ("pad", C.c_uint32), ("seed", C.c_uint64)]


ORACLE_HIT_DTYPE = np.dtype([("t", "<f8"), ("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8"), ("u", "<f8"), ("v", "<f8"),
                             ("px", "<f8"), ("py", "<f8"), ("pz", "<f8"), ("prim", "<i4"), ("front_face", "<i4")])
_lib = None


def build():
    """make -C oracle (g++ only; the product's build.py calls the same recipe)."""
    res = subprocess.run(["make", "-C", HERE], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if res.returncode != 0:
        raise RuntimeError("oracle build failed:\n" + res.stdout)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_PATH):
            build()
        h = C.CDLL(ORACLE_PATH)
        VP = C.c_void_p
        h.oracle_last_error.restype = C.c_char_p
        h.oracle_scene_create.restype = VP
        h.oracle_scene_create.argtypes = [VP, C.c_uint64]
        h.oracle_scene_make.restype = VP
        h.oracle_scene_make.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.c_char_p, VP, C.c_int32, C.c_int32, C.c_double, C.c_double, C.POINTER(Camera), C.POINTER(Params)]
        h.oracle_scene_load_json.restype = VP
        h.oracle_scene_load_json.argtypes = [C.c_char_p, C.c_uint64, C.c_uint64, C.POINTER(Camera)]
        h.oracle_scene_prim_count.restype = C.c_uint64
        h.oracle_scene_prim_count.argtypes = [VP]
        h.oracle_scene_material_count.argtypes = [VP]
        h.oracle_scene_material_info.argtypes = [VP, C.c_int, C.POINTER(C.c_int32), C.POINTER(C.c_double), C.POINTER(C.c_int32)]
        h.oracle_scene_texture_of_material.argtypes = [VP, C.c_int]
        h.oracle_scene_destroy.argtypes = [VP]
        h.oracle_hit_batch.argtypes = [VP, VP, VP, VP, C.c_uint64, C.c_double, VP]
        h.oracle_render.argtypes = [VP, VP, VP, VP, C.c_int, C.c_int, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        h.oracle_tonemap.argtypes = [VP, C.c_uint64, VP]
        h.oracle_vec3_op.argtypes = [C.c_int, VP, VP, VP]
        h.oracle_sphere_union_box.argtypes = [VP, C.c_uint32, VP]
        h.oracle_texture_eval.argtypes = [VP, C.c_int, C.c_uint64, VP, VP]
        h.oracle_scatter_eval.argtypes = [VP, C.c_uint64, VP, VP]
        h.oracle_camera_eval.argtypes = [VP, C.c_int, C.c_int, C.c_uint64, VP, VP]
        h.oracle_sppm_render.argtypes = [VP, VP, VP, VP, VP, VP, VP, VP, C.c_int]
        _lib = h
    return _lib


def _err():
    return RuntimeError("oracle: " + lib().oracle_last_error().decode())


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class OracleScene:
    """A scene of the oracle.  OracleScene.make(name) / .load_json(text) build the reference-style object graph inside the
    oracle (independent of the product); OracleScene(desc) rebuilds it from flat ABI tables (ad-hoc test scenes)."""

    def __init__(self, builder_or_desc=None, bvh_seed=7, _handle=None):
        self.camera, self.params = None, None
        if _handle is not None:
            self._h = _handle
            return
        self._keep = builder_or_desc
        desc = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        self._h = lib().oracle_scene_create(C.byref(desc), bvh_seed)
        if not self._h:
            raise _err()

    @classmethod
    def make(cls, name, seed=1, aux_path=None, earth=None, aspect=0.0, bvh_seed=7, tri_pad=0.0):
        cam, par = Camera(), Params()
        ep, ew, eh = None, 0, 0
        if earth is not None:
            earth = np.ascontiguousarray(earth, dtype=np.uint8)
            eh, ew = earth.shape[:2]
            ep = _ptr(earth)
        h = lib().oracle_scene_make(name.encode(), seed, bvh_seed, aux_path.encode() if aux_path else None, ep, ew, eh, aspect, tri_pad, C.byref(cam), C.byref(par))
        if not h:
            raise _err()
        s = cls(_handle=h)
        s.camera, s.params = cam, par
        return s

    @classmethod
    def load_json(cls, text, bvh_seed=7):
        data = text.encode() if isinstance(text, str) else bytes(text)
        cam = Camera()
        h = lib().oracle_scene_load_json(data, len(data), bvh_seed, C.byref(cam))
        if not h:
            raise _err()
        s = cls(_handle=h)
        s.camera = cam
        return s

    def __del__(self):
        try:
            if self._h:
                lib().oracle_scene_destroy(self._h)
                self._h = None
        except Exception:
            pass

    @property
    def n_prims(self):
        return int(lib().oracle_scene_prim_count(self._h))

    def materials(self):
        """[(type, param, texture type, texture index)] in construction order."""
        out = []
        for i in range(lib().oracle_scene_material_count(self._h)):
            t, p, tt = C.c_int32(), C.c_double(), C.c_int32()
            lib().oracle_scene_material_info(self._h, i, C.byref(t), C.byref(p), C.byref(tt))
            out.append((t.value, p.value, tt.value, lib().oracle_scene_texture_of_material(self._h, i)))
        return out

    def hit_batch(self, o_time, d_tmax, medium_seed=None, t_min=0.001):
        o = np.ascontiguousarray(o_time, dtype=np.float32).reshape(-1, 4)
        d = np.ascontiguousarray(d_tmax, dtype=np.float32).reshape(-1, 4)
        n = o.shape[0]
        out = np.zeros(n, ORACLE_HIT_DTYPE)
        sp = None
        if medium_seed is not None:
            medium_seed = np.ascontiguousarray(medium_seed, dtype=np.uint32)
            sp = _ptr(medium_seed)
        if lib().oracle_hit_batch(self._h, _ptr(o), _ptr(d), sp, n, t_min, _ptr(out)):
            raise _err()
        return out

    def render(self, camera, params, n_jobs=64, n_threads=8):
        """Returns (sum image [h,w,3] float64, extend_rays, shadow_rays, seconds)."""
        acc = np.zeros((params.height, params.width, 3), np.float64)
        ext, shd, sec = C.c_uint64(), C.c_uint64(), C.c_double()
        if lib().oracle_render(self._h, C.addressof(camera), C.addressof(params), _ptr(acc), n_jobs, n_threads, C.byref(ext), C.byref(shd), C.byref(sec)):
            raise _err()
        return acc, ext.value, shd.value, sec.value

    def sppm_render(self, camera, params, iterations, photons_per_iter, k_global=100, k_caustic=50, alpha=0.7, light_scale=1e6, n_threads=8):
        """SPPMIntegrator::new + capture_image.  Returns (sum image [h,w,3], per-pixel stats [h,w,10], (global, caustic) photon counts)."""
        acc = np.zeros((params.height, params.width, 3), np.float64)
        stats = np.zeros((params.height, params.width, 10), np.float64)
        counts = np.zeros(2, np.uint64)
        sp_i = np.array([iterations, photons_per_iter, k_global, k_caustic], np.int32)
        sp_d = np.array([alpha, light_scale], np.float64)
        if lib().oracle_sppm_render(self._h, C.addressof(camera), C.addressof(params), _ptr(sp_i), _ptr(sp_d), _ptr(acc), _ptr(stats), _ptr(counts), n_threads):
            raise _err()
        return acc, stats, (int(counts[0]), int(counts[1]))

    def texture_eval(self, tex, p_uv):
        p = np.ascontiguousarray(p_uv, dtype=np.float64).reshape(-1, 5)
        out = np.zeros((p.shape[0], 3), np.float64)
        if lib().oracle_texture_eval(self._h, tex, p.shape[0], _ptr(p), _ptr(out)):
            raise RuntimeError("oracle texture_eval failed")
        return out

    def scatter_eval(self, rows):
        """rows: n x 18 (material, front_face, p, n, u, v, d, sphere sample, xi, time) -> n x 12 (has_ray, dir, attenuation, emitted, interaction, -)"""
        a = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, 18)
        pad = np.zeros((a.shape[0], 20), np.float64)
        pad[:, :18] = a
        out = np.zeros((a.shape[0], 12), np.float64)
        if lib().oracle_scatter_eval(self._h, a.shape[0], _ptr(pad), _ptr(out)):
            raise RuntimeError("oracle scatter_eval failed")
        return out


def camera_eval(camera, width, height, rows):
    """rows: n x 7 (x, y, jitter x, jitter y, disk x, disk y, time draw) -> n x 7 (origin, direction, time)"""
    a = np.ascontiguousarray(rows, dtype=np.float64).reshape(-1, 7)
    out = np.zeros((a.shape[0], 7), np.float64)
    if lib().oracle_camera_eval(C.addressof(camera), width, height, a.shape[0], _ptr(a), _ptr(out)):
        raise RuntimeError("oracle camera_eval failed")
    return out


def copy_params(p, **kw):
    q = Params()
    C.memmove(C.byref(q), C.addressof(p), C.sizeof(Params))
    for k, v in kw.items():
        setattr(q, k, v)
    return q
