"""ctypes binding of oracle/liboracle.so — TEST INFRASTRUCTURE (only tests/, smoke() and bench.py's CPU legs load it)."""
import ctypes as C
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_PATH = os.path.join(ROOT, "oracle", "liboracle.so")

from rust_raytracer_b200 import _ffi as F  # struct layouts only (public ABI)

ORACLE_HIT_DTYPE = np.dtype([("t", "<f8"), ("nx", "<f8"), ("ny", "<f8"), ("nz", "<f8"), ("u", "<f8"), ("v", "<f8"),
                             ("px", "<f8"), ("py", "<f8"), ("pz", "<f8"), ("prim", "<i4"), ("front_face", "<i4")])
_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(ORACLE_PATH):
            from rust_raytracer_b200.build import build_oracle
            build_oracle()
        h = C.CDLL(ORACLE_PATH)
        h.oracle_last_error.restype = C.c_char_p
        h.oracle_scene_create.restype = C.c_void_p
        h.oracle_scene_create.argtypes = [C.POINTER(F.SceneDesc), C.c_uint64]
        h.oracle_scene_destroy.argtypes = [C.c_void_p]
        h.oracle_hit_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.c_void_p]
        h.oracle_render.argtypes = [C.c_void_p, C.POINTER(F.Camera), C.POINTER(F.Params), C.c_void_p, C.c_int, C.c_int,
                                    C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_double)]
        h.oracle_tonemap.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        h.oracle_vec3_op.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_void_p]
        h.oracle_sphere_union_box.argtypes = [C.POINTER(F.Prim), C.c_uint32, C.c_void_p]
        h.oracle_texture_eval.argtypes = [C.c_void_p, C.c_int, C.c_uint64, C.c_void_p, C.c_void_p]
        h.oracle_sppm_render.argtypes = [C.c_void_p, C.POINTER(F.Camera), C.POINTER(F.Params), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                         C.c_void_p, C.c_int]
        _lib = h
    return _lib


class OracleScene:
    def __init__(self, builder_or_desc, bvh_seed=7):
        self._keep = builder_or_desc
        desc = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        self._h = lib().oracle_scene_create(C.byref(desc), bvh_seed)
        if not self._h:
            raise RuntimeError("oracle: " + lib().oracle_last_error().decode())

    def __del__(self):
        try:
            if self._h:
                lib().oracle_scene_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def hit_batch(self, o_time, d_tmax, medium_seed=None, t_min=0.001):
        o = np.ascontiguousarray(o_time, dtype=np.float32).reshape(-1, 4)
        d = np.ascontiguousarray(d_tmax, dtype=np.float32).reshape(-1, 4)
        n = o.shape[0]
        out = np.zeros(n, ORACLE_HIT_DTYPE)
        sp = None
        if medium_seed is not None:
            medium_seed = np.ascontiguousarray(medium_seed, dtype=np.uint32)
            sp = medium_seed.ctypes.data_as(C.c_void_p)
        rc = lib().oracle_hit_batch(self._h, o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), sp, n, t_min, out.ctypes.data_as(C.c_void_p))
        if rc:
            raise RuntimeError("oracle: " + lib().oracle_last_error().decode())
        return out

    def render(self, camera, params, n_jobs=64, n_threads=8):
        """Returns (sum image [h,w,3] float64, extend_rays, shadow_rays, seconds)."""
        acc = np.zeros((params.height, params.width, 3), np.float64)
        ext, shd, sec = C.c_uint64(), C.c_uint64(), C.c_double()
        rc = lib().oracle_render(self._h, C.byref(camera), C.byref(params), acc.ctypes.data_as(C.c_void_p), n_jobs, n_threads, C.byref(ext), C.byref(shd), C.byref(sec))
        if rc:
            raise RuntimeError("oracle: " + lib().oracle_last_error().decode())
        return acc, ext.value, shd.value, sec.value

    def sppm_render(self, camera, params, iterations, photons_per_iter, k_global=100, k_caustic=50, alpha=0.7, light_scale=1e6, n_threads=8):
        """SPPMIntegrator::new + capture_image.  Returns (sum image [h,w,3], per-pixel stats [h,w,10], (global, caustic) photon counts)."""
        acc = np.zeros((params.height, params.width, 3), np.float64)
        stats = np.zeros((params.height, params.width, 10), np.float64)
        counts = np.zeros(2, np.uint64)
        sp_i = np.array([iterations, photons_per_iter, k_global, k_caustic], np.int32)
        sp_d = np.array([alpha, light_scale], np.float64)
        rc = lib().oracle_sppm_render(self._h, C.byref(camera), C.byref(params), sp_i.ctypes.data_as(C.c_void_p), sp_d.ctypes.data_as(C.c_void_p),
                                      acc.ctypes.data_as(C.c_void_p), stats.ctypes.data_as(C.c_void_p), counts.ctypes.data_as(C.c_void_p), n_threads)
        if rc:
            raise RuntimeError("oracle: " + lib().oracle_last_error().decode())
        return acc, stats, (int(counts[0]), int(counts[1]))

    def texture_eval(self, tex, p_uv):
        p = np.ascontiguousarray(p_uv, dtype=np.float64).reshape(-1, 5)
        out = np.zeros((p.shape[0], 3), np.float64)
        if lib().oracle_texture_eval(self._h, tex, p.shape[0], p.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)):
            raise RuntimeError("oracle texture_eval failed")
        return out
