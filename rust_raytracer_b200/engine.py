"""Context / Scene / render wrappers over the C ABI (numpy in, numpy out; optional raw device pointers)."""
import ctypes as C
import weakref

import numpy as np

from . import _ffi as F
from .api import B200RTError

HIT_DTYPE = np.dtype([("t", "<f4"), ("prim", "<i4"), ("nx", "<f4"), ("ny", "<f4"), ("nz", "<f4"), ("front_face", "<i4"), ("u", "<f4"), ("v", "<f4")])


class Context:
    def __init__(self, device=0):
        self._h = C.c_void_p()
        rc = F.lib().b200rt_ctx_create(device, C.byref(self._h))
        if rc:
            raise B200RTError(rc, F.lib().b200rt_global_error().decode())
        self.device = device
        # A scene should be destroyed before its context.  The weak set lets close() do that in the ordinary case; it cannot in the
        # garbage collector's (weak references to unreachable objects are cleared BEFORE their finalizers run, so a context
        # collected together with its scenes no longer sees them) — there the library's own bookkeeping takes over: it
        # orphans the scenes of a destroyed context, and b200rt_scene_destroy of an orphan only frees its tables.
        self._scenes = weakref.WeakSet()

    def _ok(self, rc):
        if rc:
            raise B200RTError(rc, F.lib().b200rt_last_error(self._h).decode())

    def close(self):
        if self._h:
            for sc in list(self._scenes):  # also at interpreter exit, where objects die in no particular order
                sc.close()
            F.lib().b200rt_ctx_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream_ptr):
        self._ok(F.lib().b200rt_ctx_set_stream(self._h, C.c_void_p(cuda_stream_ptr)))

    def device_info(self):
        sm, khz, l2 = C.c_int32(), C.c_int32(), C.c_int64()
        name = C.create_string_buffer(128)
        self._ok(F.lib().b200rt_ctx_device_info(self._h, C.byref(sm), C.byref(khz), C.byref(l2), name, 128))
        return {"sm_count": sm.value, "clock_khz": khz.value, "l2_bytes": l2.value, "name": name.value.decode()}

    def measure_fp32_peak(self):
        t, mhz = C.c_double(), C.c_double()
        self._ok(F.lib().b200rt_measure_fp32_peak(self._h, C.byref(t), C.byref(mhz)))
        return t.value, mhz.value

    def tonemap(self, accum):
        accum = np.ascontiguousarray(accum, dtype=np.float32)
        h, w = accum.shape[:2]
        out = np.empty((h, w, 3), np.uint8)
        self._ok(F.lib().b200rt_tonemap_rgb8(self._h, accum.ctypes.data_as(C.c_void_p), w, h, out.ctypes.data_as(C.c_void_p)))
        return out

    def tonemap_device(self, accum_ptr, width, height, rgb_ptr):
        self._ok(F.lib().b200rt_tonemap_rgb8_device(self._h, C.c_void_p(accum_ptr), width, height, C.c_void_p(rgb_ptr)))


class Scene:
    """World::new on the device: uploads the flat tables of a SceneBuilder (or a raw SceneDesc)."""

    def __init__(self, ctx, builder_or_desc):
        self.ctx = ctx
        self._keep = builder_or_desc
        desc = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
        self._h = C.c_void_p()
        ctx._ok(F.lib().b200rt_scene_create(ctx._h, C.byref(desc), C.byref(self._h)))
        ctx._scenes.add(self)

    def close(self):
        if self._h:
            F.lib().b200rt_scene_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def info(self):
        b, n, p = C.c_uint64(), C.c_uint32(), C.c_uint32()
        self.ctx._ok(F.lib().b200rt_scene_info(self._h, C.byref(b), C.byref(n), C.byref(p)))
        return {"h2d_bytes": b.value, "bvh_nodes": n.value, "device_prims": p.value}

    def intersect_batch(self, o_time, d_tmax, medium_seed=None, t_min=0.001):
        """World::hit over a batch: o_time (n,4) = origin + time, d_tmax (n,4) = direction + t_max."""
        o = np.ascontiguousarray(o_time, dtype=np.float32).reshape(-1, 4)
        d = np.ascontiguousarray(d_tmax, dtype=np.float32).reshape(-1, 4)
        if o.shape != d.shape:
            raise B200RTError(F.ERR_INVALID, "ragged ray batch: %s origins vs %s directions" % (o.shape, d.shape))
        n = o.shape[0]
        out = np.zeros(n, HIT_DTYPE)
        seed_ptr = None
        if medium_seed is not None:
            medium_seed = np.ascontiguousarray(medium_seed, dtype=np.uint32)
            if medium_seed.shape[0] != n:
                raise B200RTError(F.ERR_INVALID, "ragged ray batch: seeds")
            seed_ptr = medium_seed.ctypes.data_as(C.c_void_p)
        self.ctx._ok(F.lib().b200rt_intersect_batch(self._h, o.ctypes.data_as(C.c_void_p), d.ctypes.data_as(C.c_void_p), seed_ptr, n, t_min, out.ctypes.data_as(C.c_void_p)))
        return out

    def debug_eval(self, kind, camera, params, rows):
        """b200rt_debug_eval: single functions of the hot path on the device at caller-supplied inputs (F.EVAL_*).
        rows: float32 (n, 8 | 20 | 2 | 8); returns float32 (n, 4 | 16 | 16 | 12)."""
        in_w, out_w = {F.EVAL_TEXTURE: (8, 4), F.EVAL_SCATTER: (20, 16), F.EVAL_CAMERA: (2, 16), F.EVAL_SAMPLERS: (8, 12)}[kind]
        a = np.ascontiguousarray(rows, dtype=np.float32).reshape(-1, in_w)
        out = np.zeros((a.shape[0], out_w), np.float32)
        self.ctx._ok(F.lib().b200rt_debug_eval(self._h, C.byref(camera), C.byref(params), kind, a.shape[0], a.ctypes.data_as(C.c_void_p), out.ctypes.data_as(C.c_void_p)))
        return out

    def render(self, camera, params):
        """capture_image into a fresh host buffer: returns (accum[h,w,4] float32 sums + count, Stats)."""
        accum = np.zeros((params.height, params.width, 4), np.float32)
        st = F.Stats()
        self.ctx._ok(F.lib().b200rt_render(self._h, C.byref(camera), C.byref(params), accum.ctypes.data_as(C.c_void_p), C.byref(st)))
        return accum, st

    def render_sppm(self, camera, params, iterations=50, photons_per_iter=500000, k_global=100, k_caustic=50, alpha=0.7, light_scale=1e6,
                    want_pixel_stats=False):
        """SPPMIntegrator::new + capture_image (photon_mapper.rs:139-236, main.rs:51-53): returns (accum[h,w,4], SppmStats) or,
        with want_pixel_stats, (accum, per-pixel stats [h,w,10], SppmStats)."""
        accum = np.zeros((params.height, params.width, 4), np.float32)
        px = np.zeros((params.height, params.width, 10), np.float32) if want_pixel_stats else None
        sp = F.SppmParams(iterations, photons_per_iter, k_global, k_caustic, alpha, light_scale)
        st = F.SppmStats()
        self.ctx._ok(F.lib().b200rt_render_sppm(self._h, C.byref(camera), C.byref(params), C.byref(sp), accum.ctypes.data_as(C.c_void_p),
                                                px.ctypes.data_as(C.c_void_p) if want_pixel_stats else None, C.byref(st)))
        return (accum, px, st) if want_pixel_stats else (accum, st)

    def render_device_static(self, camera, params, accum_ptr):
        """b200rt_render_device_static: the same render with the queues in the library's static device array (no cudaMalloc)."""
        st = F.Stats()
        self.ctx._ok(F.lib().b200rt_render_device_static(self._h, C.byref(camera), C.byref(params), C.c_void_p(accum_ptr), C.byref(st)))
        return st

    def render_device(self, camera, params, accum_ptr):
        st = F.Stats()
        self.ctx._ok(F.lib().b200rt_render_device(self._h, C.byref(camera), C.byref(params), C.c_void_p(accum_ptr), C.byref(st)))
        return st


def render_sppm_multi(scenes, camera, params, iterations=50, photons_per_iter=500000, k_global=100, k_caustic=50, alpha=0.7, light_scale=1e6,
                      want_pixel_stats=False):
    """b200rt_render_sppm_multi: the SPPM integrator over the devices of `scenes` (photons sharded by id, the map all-gathered
    inside the grid build over peer memory, pixel statistics sharded by pixel).  Returns like Scene.render_sppm."""
    n = len(scenes)
    handles = (C.c_void_p * n)(*[s._h for s in scenes])
    accum = np.zeros((params.height, params.width, 4), np.float32)
    px = np.zeros((params.height, params.width, 10), np.float32) if want_pixel_stats else None
    sp = F.SppmParams(iterations, photons_per_iter, k_global, k_caustic, alpha, light_scale)
    st = F.SppmStats()
    scenes[0].ctx._ok(F.lib().b200rt_render_sppm_multi(n, handles, C.byref(camera), C.byref(params), C.byref(sp), accum.ctypes.data_as(C.c_void_p),
                                                       px.ctypes.data_as(C.c_void_p) if want_pixel_stats else None, C.byref(st)))
    return (accum, px, st) if want_pixel_stats else (accum, st)


def compile_stats(builder_or_desc):
    """Host only: the wide BVH scene_create would build — {"bvh_nodes", "bvh4_depth", "stack_needed", "stack_capacity"}."""
    desc = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
    v = [C.c_uint32() for _ in range(4)]
    rc = F.lib().b200rt_scene_compile_stats(C.byref(desc), *[C.byref(x) for x in v])
    if rc:
        raise B200RTError(rc, "scene_compile_stats")
    return dict(zip(("bvh_nodes", "bvh4_depth", "stack_needed", "stack_capacity"), [x.value for x in v]))


def bvh_quality(builder_or_desc):
    """Host only: the surface-area estimate of the BVH scene_create would build and the scene's dense box, if any —
    {"node_steps", "prim_tests", "dense_box": None | (lo[3], hi[3])} (b200rt_scene_bvh_quality)."""
    desc = builder_or_desc.desc() if hasattr(builder_or_desc, "desc") else builder_or_desc
    ns, pt, has = C.c_float(), C.c_float(), C.c_int32()
    lo, hi = (C.c_float * 3)(), (C.c_float * 3)()
    rc = F.lib().b200rt_scene_bvh_quality(C.byref(desc), C.byref(ns), C.byref(pt), C.byref(has), lo, hi)
    if rc:
        raise B200RTError(rc, "scene_bvh_quality")
    return {"node_steps": ns.value, "prim_tests": pt.value, "dense_box": (tuple(lo), tuple(hi)) if has.value else None}


def render_multi(scenes, camera, params):
    """capture_image sharded by sample index over the devices of `scenes` (the same scene created on one Context per GPU) with
    one NCCL / peer-memory sum-reduce onto scenes[0]'s device (b200rt_render_multi).
    Returns (accum[h,w,4], [Stats per device], MultiStats)."""
    n = len(scenes)
    handles = (C.c_void_p * n)(*[s._h for s in scenes])
    accum = np.zeros((params.height, params.width, 4), np.float32)
    st = (F.Stats * n)()
    ms = F.MultiStats()
    scenes[0].ctx._ok(F.lib().b200rt_render_multi(n, handles, C.byref(camera), C.byref(params), accum.ctypes.data_as(C.c_void_p), st, C.byref(ms)))
    return accum, list(st), ms


def make_params(width, height, spp, integrator=F.INTEGRATOR_PATH, pipeline=F.PIPELINE_WAVEFRONT, max_depth=50, t_min=0.001,
                background=(0.0, 0.0, 0.0), background_mode=F.BACKGROUND_CONSTANT, seed=1, spp_begin=0, flags=0):
    p = F.Params()
    p.width, p.height, p.spp_begin, p.spp_count, p.max_depth = width, height, spp_begin, spp, max_depth
    p.integrator, p.pipeline, p.background_mode, p.t_min = integrator, pipeline, background_mode, t_min
    p.background[0], p.background[1], p.background[2] = background
    p.flags, p.seed = flags, seed
    return p


def copy_params(p, **kw):
    q = F.Params()
    C.memmove(C.byref(q), C.byref(p), C.sizeof(F.Params))
    for k, v in kw.items():
        setattr(q, k, v)
    return q
