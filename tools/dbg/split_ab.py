"""A/B of the two-ended bins by the scene's dense box (WfBuffers::split_box) on one scene, in one process:
the scene is built once, B200RT_WF_SPLIT is switched between calls (render_wavefront reads it per call).

  python tools/dbg/split_ab.py [scene=mesh] [w h spp integrator] [rounds] [modes=0,2] [variable=B200RT_WF_SPLIT]
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "mesh"
w, h, spp = (int(x) for x in sys.argv[2:5]) if len(sys.argv) > 4 else (1920, 1080, 256)
integ = {"path": F.INTEGRATOR_PATH, "mixture": F.INTEGRATOR_MIXTURE, "nee": F.INTEGRATOR_NEE}[sys.argv[5] if len(sys.argv) > 5 else "mixture"]
rounds = int(sys.argv[6]) if len(sys.argv) > 6 else 3
modes = (sys.argv[7] if len(sys.argv) > 7 else "0,2").split(",")
var = sys.argv[8] if len(sys.argv) > 8 else "B200RT_WF_SPLIT"  # the variable the modes are values of
ctx = rt.Context(0)
sb = scene_builder(name)
scene = rt.Scene(ctx, sb)
p = rt.copy_params(sb.params, width=w, height=h, spp_count=spp, integrator=integ)
scene.render(sb.camera, p)  # warm-up: allocates the pools
imgs = {}
for r in range(rounds):
    for mode in modes:
        os.environ[var] = mode
        acc, st = scene.render(sb.camera, rt.copy_params(p, spp_begin=spp))
        imgs[mode] = acc[..., :3].astype(np.float64) / spp
        print(json.dumps({"scene": name, "var": var, "split": mode, "round": r, "ms": round(st.ms_total, 3), "mpaths_s": round(st.paths / st.ms_total / 1e3, 1),
                          "mrays_s": round((st.extend_rays + st.shadow_rays) / st.ms_total / 1e3, 1), "extend_rays": st.extend_rays, "shadow_rays": st.shadow_rays,
                          "ms_extend_sum": round(st.ms_extend_sum, 2), "ms_shade_sum": round(st.ms_shade_sum, 2), "iterations": st.wavefront_iterations, "pipelines": st.pipelines}), flush=True)
a, b = imgs[modes[0]], imgs[modes[-1]]
print(json.dumps({"scene": name, "image_max_abs_diff": float(np.abs(a - b).max()), "image_mean": a.mean(axis=(0, 1)).tolist(), "image_mean_diff": (a.mean(axis=(0, 1)) - b.mean(axis=(0, 1))).tolist()}))
