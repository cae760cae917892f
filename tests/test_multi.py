"""b200rt_render_multi: capture_image sharded by sample index over the GPUs of one node with
one NCCL (or peer-memory) sum-reduce, inside the C ABI.  The 2-GPU cases need `gpurun --gpus 2`; on a one-GPU box
they skip and the one-device identities still run."""
import ctypes as C
import os

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from scenes_util import scene_builder


def test_multi_split_matches_sharding_py():
    """The C split of the sample range (api_multi.cuh multi_spp_range) is restated here and must equal sharding.spp_range."""
    from rust_raytracer_b200.sharding import spp_range
    for world in (1, 2, 3, 4, 8):
        for spp in (0, 1, 7, 8, 100, 10000):
            for begin in (0, 5):
                got = []
                for rank in range(world):
                    base, rem = divmod(spp, world)
                    got.append((begin + rank * base + min(rank, rem), base + (1 if rank < rem else 0)))
                assert got == [spp_range(r, world, spp, begin) for r in range(world)]
                assert sum(c for _, c in got) == spp


def _n_devices():
    n = C.c_int(0)
    F.lib().b200rt_device_count(C.byref(n))
    return n.value


@pytest.mark.gpu
def test_render_multi_one_device_is_render(ctx):
    sb = scene_builder("cornell_book3")
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=48, height=40, spp_count=6, spp_begin=3)
    want, st1 = scene.render(sb.camera, p)
    got, st, ms = rt.render_multi([scene], sb.camera, p)
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)  # per-pixel float atomics: the order of the adds is free
    assert ms.n_dev == 1 and ms.reduce_path == F.REDUCE_NONE and ms.paths == 48 * 40 * 6 == st[0].paths
    assert ms.extend_rays == st1.extend_rays


@pytest.mark.gpu
def test_render_multi_argument_errors(ctx):
    sb = scene_builder("cornell_book3")
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=16, height=16, spp_count=2)
    with pytest.raises(rt.B200RTError) as e:
        rt.render_multi([scene, scene], sb.camera, p)  # two shards on one context (a context is not thread-safe)
    assert e.value.status == F.ERR_INVALID and "own context" in str(e.value)
    with pytest.raises(rt.B200RTError):
        rt.render_multi([scene], sb.camera, rt.copy_params(p, width=0))
    ctx_b = rt.Context(0)
    old = os.environ.get("B200RT_MULTI_REDUCE")
    os.environ["B200RT_MULTI_REDUCE"] = "nccl"
    try:
        with pytest.raises(rt.B200RTError) as e:  # NCCL cannot rank one device twice
            rt.render_multi([scene, rt.Scene(ctx_b, sb)], sb.camera, p)
        assert "share a device" in str(e.value)
    finally:
        os.environ.pop("B200RT_MULTI_REDUCE") if old is None else os.environ.__setitem__("B200RT_MULTI_REDUCE", old)


@pytest.mark.gpu
@pytest.mark.parametrize("n_ctx", [2, 3])
def test_render_multi_on_one_device_through_the_peer_reduce(ctx, n_ctx):
    """The whole multi-device entry point on ONE GPU (the driver's box): n contexts on device 0, one host thread each, disjoint
    contiguous sample ranges, then k_reduce_peers sums the per-context buffers (plain local memory here, NVLink peers on a
    multi-GPU box) — against the same sample ranges rendered one by one and added on the host."""
    sb = scene_builder("book2")
    ctxs = [ctx] + [rt.Context(0) for _ in range(n_ctx - 1)]
    scenes = [rt.Scene(c, sb) for c in ctxs]
    p = rt.copy_params(sb.params, width=96, height=80, spp_count=11, spp_begin=4)  # 11 = 6 + 5 or 4 + 4 + 3
    from rust_raytracer_b200.sharding import spp_range
    want = np.zeros((80, 96, 4), np.float32)
    rays = 0
    for r in range(n_ctx):
        b, c = spp_range(r, n_ctx, 11, 4)
        img, st = scenes[0].render(sb.camera, rt.copy_params(p, spp_begin=b, spp_count=c))
        want += img
        rays += st.extend_rays
    got, st, ms = rt.render_multi(scenes, sb.camera, p)
    assert ms.n_dev == n_ctx and ms.reduce_path == F.REDUCE_PEER and ms.paths == 96 * 80 * 11 and ms.extend_rays == rays
    assert [s.paths for s in st] == [96 * 80 * spp_range(r, n_ctx, 11, 4)[1] for r in range(n_ctx)]
    np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)
    assert np.all(got[..., 3] == 11)
    full, _ = scenes[0].render(sb.camera, p)
    np.testing.assert_allclose(got, full, rtol=1e-4, atol=1e-4)
    print("render_multi, %d contexts on one device: render %.2f ms, peer reduce %.3f ms" % (n_ctx, ms.ms_render_max, ms.ms_reduce))


@pytest.mark.gpu
@pytest.mark.parametrize("path", ["nccl", "peer"])
def test_render_multi_two_devices(ctx, path):
    if _n_devices() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    sb = scene_builder("book2")
    ctx1 = rt.Context(1)
    s0, s1 = rt.Scene(ctx, sb), rt.Scene(ctx1, sb)
    p = rt.copy_params(sb.params, width=96, height=96, spp_count=9)  # odd: 5 + 4
    full, _ = s0.render(sb.camera, p)
    lo, _ = s0.render(sb.camera, rt.copy_params(p, spp_begin=0, spp_count=5))
    hi, _ = s0.render(sb.camera, rt.copy_params(p, spp_begin=5, spp_count=4))
    old = os.environ.get("B200RT_MULTI_REDUCE")
    os.environ["B200RT_MULTI_REDUCE"] = path
    try:
        got, st, ms = rt.render_multi([s0, s1], sb.camera, p)
    finally:
        if old is None:
            del os.environ["B200RT_MULTI_REDUCE"]
        else:
            os.environ["B200RT_MULTI_REDUCE"] = old
    assert ms.n_dev == 2 and ms.reduce_path == (F.REDUCE_NCCL if path == "nccl" else F.REDUCE_PEER)
    assert st[0].paths == 96 * 96 * 5 and st[1].paths == 96 * 96 * 4 and ms.paths == 96 * 96 * 9
    # device 1 rendered what device 0 renders for the same sample range (up to the order of the per-pixel float adds)
    np.testing.assert_allclose(got, lo + hi, rtol=1e-4, atol=1e-4)
    np.testing.assert_allclose(got, full, rtol=1e-4, atol=1e-4)
    assert np.all(got[..., 3] == 9)
    print("render_multi[%s]: render %.2f ms, reduce %.3f ms, wall %.1f ms" % (path, ms.ms_render_max, ms.ms_reduce, ms.ms_wall))


@pytest.mark.gpu
@pytest.mark.parametrize("integrator", ["path", "nee"])
def test_concurrent_pipelines_render_the_same_paths(ctx, integrator):
    """render_wavefront splits the sample range over concurrent pipelines (own pool / counters / stream, one accumulation
    buffer): the paths are the same ones (Philox is keyed on the global sample index), so the ray counts are identical
    and the image equal up to the order of the per-pixel float adds."""
    sb = scene_builder("cornell_book3")
    scene = rt.Scene(ctx, sb)
    kw = dict(integrator=F.INTEGRATOR_NEE) if integrator == "nee" else {}
    p = rt.copy_params(sb.params, width=64, height=48, spp_count=7, spp_begin=2, **kw)
    out = {}
    old = {k: os.environ.get(k) for k in ("B200RT_WF_PIPES", "B200RT_WF_PIPE_MIN_LOG2")}
    try:
        os.environ["B200RT_WF_PIPE_MIN_LOG2"] = "0"
        for n in (1, 2, 3, 4):
            os.environ["B200RT_WF_PIPES"] = str(n)
            out[n] = scene.render(sb.camera, p)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v
    a1, s1 = out[1]
    assert s1.paths == 64 * 48 * 7
    for n in (2, 3, 4):
        a, s = out[n]
        assert (s.paths, s.extend_rays, s.shadow_rays) == (s1.paths, s1.extend_rays, s1.shadow_rays), n
        assert s.pipelines == n and s.wavefront_iterations >= s1.wavefront_iterations  # iterations are counted per pipeline (batches of 4, then the tail kernel)
        np.testing.assert_allclose(a, a1, rtol=1e-4, atol=1e-4)
        assert np.all(a[..., 3] == 7)


@pytest.mark.gpu
def test_context_destroyed_before_its_scene_orphans_the_scene(ctx):
    """A garbage collector may finalize a context before its scenes (weak references to unreachable objects are cleared before
    their finalizers run).  The library orphans the scenes of a destroyed context: entry points answer ERR_INVALID instead of
    following a dangling pointer, b200rt_scene_destroy still frees the tables, and no CUDA error is left behind for the next call."""
    import ctypes as C
    sb = scene_builder("cornell_ref")
    c2 = rt.Context(0)
    sc = rt.Scene(c2, sb)
    h_ctx, c2._h = c2._h, None  # destroy the context behind the wrapper's back, as the collector's order would
    assert F.lib().b200rt_ctx_destroy(h_ctx) == 0
    acc = np.zeros((16, 16, 4), np.float32)
    st = F.Stats()
    p = rt.copy_params(sb.params, width=16, height=16, spp_count=1)
    rc = F.lib().b200rt_render(sc._h, C.byref(sb.camera), C.byref(p), acc.ctypes.data_as(C.c_void_p), C.byref(st))
    assert rc == F.ERR_INVALID and b"context has been destroyed" in F.lib().b200rt_global_error()
    h_sc, sc._h = sc._h, None
    assert F.lib().b200rt_scene_destroy(h_sc) == 0
    # the session context is untouched and sees no stale error
    a, _ = rt.Scene(ctx, sb).render(sb.camera, p)
    assert np.isfinite(a).all() and a[..., 3].min() == 1


@pytest.mark.gpu
def test_garbage_collected_context_and_scene_cycle(ctx):
    """The collector's order, for real: a context and its scene die together inside a reference cycle.  Weak references to
    unreachable objects are cleared before their finalizers run, so Context.close() no longer finds the scene; whichever
    finalizer runs first, nothing may crash and no CUDA error may be left for the next call."""
    import gc
    sb = scene_builder("cornell_ref")
    for _ in range(3):
        c2 = rt.Context(0)
        sc = rt.Scene(c2, sb)
        sc.cycle = [sc, c2]  # unreachable only through the cycle collector
        c2.cycle = sc
        del c2, sc
        gc.collect()
    p = rt.copy_params(sb.params, width=16, height=16, spp_count=1)
    a, _ = rt.Scene(ctx, sb).render(sb.camera, p)
    assert np.isfinite(a).all() and a[..., 3].min() == 1
