"""b200rt_render_device_static (no counterpart in the reference): capture_image
with the wavefront queues in the library's file-scope static device array instead of a cudaMalloc'ed pool.  Called through the
C ABI on two calls in a row and checked against the existing entry point, b200rt_render_device: the Philox keys are the same, so
both trace the same paths (equal ray counts) and the images agree up to the order of the per-pixel float adds (the tolerance
the other pipeline-equivalence tests use).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

import rust_raytracer_b200 as rt
from rust_raytracer_b200 import _ffi as F
from scenes_util import scene_builder

pytestmark = pytest.mark.gpu


def _device_image(torch, scene, sb, p, static):
    acc = torch.zeros((p.height, p.width, 4), dtype=torch.float32, device="cuda:0")
    torch.cuda.synchronize()
    st = (scene.render_device_static if static else scene.render_device)(sb.camera, p, acc.data_ptr())
    torch.cuda.synchronize()
    return acc.cpu().numpy(), st


@pytest.mark.parametrize("name,integrator", [("book2", F.INTEGRATOR_PATH), ("cornell_book3", F.INTEGRATOR_NEE)])
def test_static_queue_render_equals_render_device(name, integrator):
    torch = pytest.importorskip("torch")
    ctx = rt.Context(0)  # own context: passes alone in a fresh process and after the other tests alike
    sb = scene_builder(name)
    scene = rt.Scene(ctx, sb)
    # 96 x 80 x 24 = 184,320 paths: about three times the 65,536 queue positions of the static block, so the queues refill
    for call, begin in enumerate((0, 24)):  # two calls in a row: the second one meets the first one's leftovers in the array
        p = rt.copy_params(sb.params, width=96, height=80, spp_count=24, spp_begin=begin, integrator=integrator)
        want, st_w = _device_image(torch, scene, sb, p, static=False)
        got, st_g = _device_image(torch, scene, sb, p, static=True)
        assert st_g.paths == st_w.paths == 96 * 80 * 24
        assert (st_g.extend_rays, st_g.shadow_rays) == (st_w.extend_rays, st_w.shadow_rays), call
        assert st_g.pipelines == 1 and st_g.wavefront_iterations >= 3
        np.testing.assert_allclose(got, want, rtol=1e-4, atol=1e-4)
        assert np.all(got[..., 3] == 24) and got[..., :3].sum() > 0
    scene.close()
    ctx.close()


def test_static_queue_render_rejects_oversize_and_megakernel():
    torch = pytest.importorskip("torch")
    ctx = rt.Context(0)
    sb = scene_builder("cornell_book3")
    scene = rt.Scene(ctx, sb)
    acc = torch.zeros((4096, 4096, 4), dtype=torch.float32, device="cuda:0")
    with pytest.raises(rt.B200RTError) as e:  # 4096 x 4096 x 2 = 2^25 paths > 2^24
        scene.render_device_static(sb.camera, rt.copy_params(sb.params, width=4096, height=4096, spp_count=2), acc.data_ptr())
    assert e.value.status == F.ERR_INVALID and "exceed the limit" in str(e.value)
    with pytest.raises(rt.B200RTError) as e:
        scene.render_device_static(sb.camera, rt.copy_params(sb.params, width=16, height=16, spp_count=1, pipeline=F.PIPELINE_MEGAKERNEL), acc.data_ptr())
    assert e.value.status == F.ERR_INVALID and "wavefront" in str(e.value)
    # at the limit it runs: 4096 x 4096 x 1 = 2^24 paths
    st = scene.render_device_static(sb.camera, rt.copy_params(sb.params, width=4096, height=4096, spp_count=1), acc.data_ptr())
    torch.cuda.synchronize()
    assert st.paths == 1 << 24
    scene.close()
    ctx.close()
