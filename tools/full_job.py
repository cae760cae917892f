"""BASELINE.json's target job through the C ABI: book-2 final scene, 800x800, 10,000 spp, on all GPUs of the node in ONE
b200rt_render_multi call (sample ranges per device + one peer-memory / NCCL reduce), then tonemap + PNG.  Prints one JSON
line with the device time (slowest GPU's CUDA-event render time + the reduce), the host wall time of the call, and a
consistency check against an independent one-GPU render (different samples): per-channel image means.
--scene mesh runs BASELINE configs[4] the same way: the 317,952-triangle mesh in the Cornell box, 1920x1080, 4096 spp, MixturePDF.
usage: python tools/full_job.py [--scene book2|mesh] [--spp N] [--gpus N] [--check-spp 1024] [--out gpurun_out/book2_full.png]"""
import argparse
import ctypes as C
import json
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

ap = argparse.ArgumentParser()
ap.add_argument("--scene", default="book2", choices=["book2", "mesh"])
ap.add_argument("--spp", type=int, default=0)
ap.add_argument("--gpus", type=int, default=0)
ap.add_argument("--check-spp", type=int, default=1024)
ap.add_argument("--out", default="")
a = ap.parse_args()
n = C.c_int(0)
F.lib().b200rt_device_count(C.byref(n))
gpus = a.gpus or n.value
sb = scene_builder(a.scene)
W, H = (800, 800) if a.scene == "book2" else (1920, 1080)
a.spp = a.spp or (10000 if a.scene == "book2" else 4096)
ctxs = [rt.Context(g) for g in range(gpus)]
scenes = [rt.Scene(c, sb) for c in ctxs]
p = rt.copy_params(sb.params, width=W, height=H, spp_count=a.spp, seed=1)
rt.render_multi(scenes, sb.camera, rt.copy_params(p, spp_count=8 * gpus, spp_begin=1 << 20))  # warm-up: pools, peer mappings
t0 = time.perf_counter()
acc, st, ms = rt.render_multi(scenes, sb.camera, p)
wall = time.perf_counter() - t0
paths = W * H * a.spp
dev_ms = ms.ms_render_max + ms.ms_reduce
what = "book-2 final scene" if a.scene == "book2" else "317,952-triangle mesh in the Cornell box (MixturePDF)"
out = {"job": "%s %dx%d, %d spp, depth 50, %d x B200, one b200rt_render_multi call" % (what, W, H, a.spp, gpus),
       "paths": paths, "rays": int(ms.extend_rays + ms.shadow_rays),
       "device_ms": dev_ms, "ms_render_max": ms.ms_render_max, "ms_render_per_gpu": [round(s.ms_total, 1) for s in st], "ms_reduce": ms.ms_reduce,
       "reduce_path": {0: "none", 1: "nccl", 2: "peer"}[ms.reduce_path], "wall_ms": wall * 1e3,
       "Mpaths_per_s_device": paths / dev_ms / 1e3, "Mrays_per_s_device": (ms.extend_rays + ms.shadow_rays) / dev_ms / 1e3,
       "Mpaths_per_s_wall": paths / wall / 1e6}
img = acc[..., :3].astype(np.float64) / acc[..., 3:4]
assert np.all(acc[..., 3] == a.spp)
out["image_mean"] = [float(x) for x in img.reshape(-1, 3).mean(axis=0)]
if a.check_spp > 0:
    acc1, st1 = scenes[0].render(sb.camera, rt.copy_params(p, spp_count=a.check_spp, spp_begin=a.spp + 7))  # samples the job did not use
    img1 = acc1[..., :3].astype(np.float64) / acc1[..., 3:4]
    m1 = img1.reshape(-1, 3).mean(axis=0)
    out["check"] = {"what": "independent one-GPU render, %d other samples per pixel" % a.check_spp, "image_mean": [float(x) for x in m1],
                    "rel_diff_of_means": [float(abs(x - y) / y) for x, y in zip(out["image_mean"], m1)],
                    "one_gpu_Mpaths_per_s": W * H * a.check_spp / st1.ms_total / 1e3}
    # 16x16 tile means: the 10k-spp job against the check render (whose noise dominates)
    t = lambda im: im[:H // 16 * 16].reshape(H // 16, 16, W // 16, 16, 3).mean(axis=(1, 3))  # noqa: E731
    out["check"]["tile_rms_rel"] = float(np.sqrt(np.mean((t(img) - t(img1)) ** 2)) / np.mean(t(img1)))
if a.out:
    rgb = ctxs[0].tonemap(acc)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    F.lib().b200rt_write_png(a.out.encode(), rgb.ctypes.data_as(C.c_void_p), W, H)
    out["png"] = a.out
print(json.dumps(out), flush=True)
for sc in scenes:
    sc.close()
for c in ctxs:
    c.close()
