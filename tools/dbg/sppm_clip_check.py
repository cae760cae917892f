"""The radius gather walks, per grid row, only the cells the pixel's SPHERE reaches (sppm.cuh, SPPM_GATHER_CLIP).  A clipping bug
would drop photons silently (the parity tests are statistical).  This renders the same SPPM frame with the shipped library and
with a -DSPPM_GATHER_CLIP=0 build (whole box per row) and compares the per-pixel photon COUNTS (integers, order-independent),
radius^2 and flux.   python tools/dbg/sppm_clip_check.py   (builds tools/dbg/libb200rt_noclip.so if missing)"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

if len(sys.argv) > 1 and sys.argv[1] == "--child":
    import numpy as np
    import rust_raytracer_b200 as rt
    from scenes_util import scene_builder
    sb = scene_builder("cornell_ref")
    sc = rt.Scene(rt.Context(0), sb)
    acc, px, st = sc.render_sppm(sb.camera, rt.copy_params(sb.params, width=200, height=200, spp_count=2, seed=3), iterations=8, photons_per_iter=100000, want_pixel_stats=True)
    np.save(sys.argv[2], px)
    print(json.dumps({"global": int(st.global_photons), "caustic": int(st.caustic_photons)}))
    sys.exit(0)

import numpy as np  # noqa: E402
noclip = os.path.join(ROOT, "tools", "dbg", "libb200rt_noclip.so")
if not os.path.exists(noclip):
    from rust_raytracer_b200.build import build_variant
    build_variant("noclip", ["SPPM_GATHER_CLIP=0"])
out = {}
for name, lib in (("clip", ""), ("noclip", noclip)):
    env = dict(os.environ)
    if lib:
        env["B200RT_LIB"] = lib
    f = os.path.join(ROOT, "gpurun_out", "clip_%s.npy" % name)
    os.makedirs(os.path.dirname(f), exist_ok=True)
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", f], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True)
    if r.returncode:
        print(name, "FAILED", r.stderr[-1500:])
        sys.exit(1)
    out[name] = np.load(f)
a, b = out["clip"], out["noclip"]
res = {"pixels": int(a.shape[0] * a.shape[1]),
       "photon_count_mismatches_global": int((a[..., 4] != b[..., 4]).sum()), "photon_count_mismatches_caustic": int((a[..., 9] != b[..., 9]).sum()),
       "radius2_max_rel_diff": float(np.max(np.abs(a[..., [3, 8]] - b[..., [3, 8]]) / np.maximum(np.abs(b[..., [3, 8]]), 1e-30))),
       "flux_max_rel_diff": float(np.max(np.abs(a[..., [0, 1, 2, 5, 6, 7]] - b[..., [0, 1, 2, 5, 6, 7]]) / (np.abs(b[..., [0, 1, 2, 5, 6, 7]]) + 1e-3 * np.abs(b[..., [0, 1, 2, 5, 6, 7]]).max())))}
print(json.dumps(res))
sys.exit(0 if res["photon_count_mismatches_global"] == 0 and res["photon_count_mismatches_caustic"] == 0 else 2)
