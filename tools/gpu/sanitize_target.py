"""Small workload for compute-sanitizer (memcheck / racecheck): every kernel family once, sized for a 50x slowdown.
  compute-sanitizer --tool memcheck  python tools/gpu/sanitize_target.py
  compute-sanitizer --tool racecheck python tools/gpu/sanitize_target.py
Covers: k_intersect_batch, the wavefront pipeline with TWO concurrent pipelines adding into one accumulation buffer (PATH on
book-2: out-of-BVH tail + warp votes; NEE on the Cornell box: shadow queue + warp-aggregated appends, and the object-space arm
of its rotated boxes), the megakernel, SPPM (photon pass overlapped with the grid build / eye pass on a second stream), tonemap."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ["B200RT_WF_PIPE_MIN_LOG2"] = "0"   # two pipelines even on a tiny render
os.environ["B200RT_WF_POOL_LOG2"] = "14"      # small pools: several wavefront iterations with refill
import numpy as np  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder, ray_batch  # noqa: E402

ctx = rt.Context(0)
sb = scene_builder("book2")
scene = rt.Scene(ctx, sb)
o, d, seeds = ray_batch(sb, 3000, np.random.default_rng(0), cam_res=(24, 16))
h = scene.intersect_batch(o, d, seeds)
print("intersect_batch", len(h), int((h["prim"] >= 0).sum()))
p = rt.copy_params(sb.params, width=48, height=32, spp_count=6)
a, st = scene.render(sb.camera, p)
print("wavefront path", st.paths, st.extend_rays, st.wavefront_iterations, st.pipelines)
b, st2 = scene.render(sb.camera, rt.copy_params(p, pipeline=F.PIPELINE_MEGAKERNEL))
print("megakernel", st2.extend_rays, float(np.abs(a - b).max()))
sc3 = scene_builder("cornell_book3")
s3 = rt.Scene(ctx, sc3)
for integ in (F.INTEGRATOR_NEE, F.INTEGRATOR_MIXTURE):
    c, st3 = s3.render(sc3.camera, rt.copy_params(sc3.params, width=40, height=40, spp_count=4, integrator=integ))
    print("cornell_book3 integrator", integ, st3.extend_rays, st3.shadow_rays, st3.pipelines)
scr = scene_builder("cornell_ref")
sr = rt.Scene(ctx, scr)
acc, sst = sr.render_sppm(scr.camera, rt.copy_params(scr.params, width=24, height=24, spp_count=2), iterations=3, photons_per_iter=4000)
print("sppm", sst.global_photons, sst.caustic_photons, sst.photon_rays)
print("tonemap", ctx.tonemap(acc).shape)
print("SANITIZE_TARGET_DONE")
