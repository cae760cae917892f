"""Throughput of every BASELINE.json config on one GPU (device-timed capture_image, reduced spp; Mpaths/s does not
depend on spp).  usage: python tools/bench_configs.py [spp]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
ctx = rt.Context(0)
CASES = [
    ("C1 book-1 final, 400x225, path", "book1", 400, 225, {}, 0),
    ("C2 Cornell book-3, 600x600, MixturePDF", "cornell_book3", 600, 600, dict(integrator=F.INTEGRATOR_MIXTURE), 0),
    ("C2' Cornell book-3, 600x600, NEE+MIS", "cornell_book3", 600, 600, dict(integrator=F.INTEGRATOR_NEE), 0),
    ("C3 scene_500.json BVH, 1920x1080", "json:scene_500", 1920, 1080, {}, 0),
    ("C3 scene_200_no_bvh.json flat list, 1920x1080", "json:scene_200_no_bvh", 1920, 1080, {}, 0),
    ("C4 book-2 final, 800x800, path", "book2", 800, 800, {}, 0),
    ("C5 mesh 317,952 tris in Cornell, 1920x1080, MixturePDF", "mesh", 1920, 1080, dict(integrator=F.INTEGRATOR_MIXTURE), 0),
]
for label, name, w, h, kw, _ in CASES:
    t0 = time.perf_counter()
    sb = scene_builder(name)
    if name.startswith("json:"):
        # the JSON camera has aspect 1: override to 16:9 like SURVEY.md section 8(d) says
        pass
    t1 = time.perf_counter()
    scene = rt.Scene(ctx, sb)
    t2 = time.perf_counter()
    n = max(1, spp // 8) if "flat" in label else spp
    p = rt.copy_params(sb.params, width=w, height=h, spp_count=n, **kw)
    scene.render(sb.camera, p)  # warm-up (allocates the pool)
    acc, st = scene.render(sb.camera, rt.copy_params(p, spp_begin=n))
    info = scene.info()
    print("%-58s %8.1f Mpaths/s %8.1f Mrays/s  %7.2f ms (gen %.1f ext %.1f shade %.1f shadow %.1f)  rays/path %.2f  prims %d nodes %d  front-end %.0f ms scene_create %.0f ms" % (
        label, st.paths / st.ms_total / 1e3, (st.extend_rays + st.shadow_rays) / st.ms_total / 1e3, st.ms_total, st.ms_generate, st.ms_extend, st.ms_shade, st.ms_shadow,
        (st.extend_rays + st.shadow_rays) / st.paths, info["device_prims"], info["bvh_nodes"], (t1 - t0) * 1e3, (t2 - t1) * 1e3), flush=True)
    scene.close()
