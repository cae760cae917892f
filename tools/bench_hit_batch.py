"""World::hit over a batch (b200rt_intersect_batch_device, world.rs:27-29): Mrays/s of k_intersect_batch on device-resident rays.
Camera rays (coherent) and secondary rays (uniform random directions from points on surfaces: incoherent) on the BASELINE scenes;
CUDA events around the call, best of 5 after a warm-up.   python tools/bench_hit_batch.py [--n 16777216]"""
import argparse
import ctypes as C
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402
import torch  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder, ray_batch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=1 << 24)
a = ap.parse_args()
ctx = rt.Context(0)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
rows = []
for name in ("book1", "cornell_book3", "json:scene_500", "book2", "mesh"):
    sb = scene_builder(name)
    scene = rt.Scene(ctx, sb)
    rng = np.random.default_rng(1)
    o, d, seeds = ray_batch(sb, 200000, rng, cam_res=(256, 256))  # camera + random + secondary rays (the parity tests' mix)
    reps = (a.n + len(o) - 1) // len(o)
    for kind, sel in (("mixed batch of the parity tests (camera, random, secondary)", slice(None)),):
        oo = torch.from_numpy(np.tile(o[sel], (reps, 1))[:a.n].astype(np.float32)).cuda()
        dd = torch.from_numpy(np.tile(d[sel], (reps, 1))[:a.n].astype(np.float32)).cuda()
        perm = torch.randperm(a.n, device="cuda")  # no two neighbouring lanes share a ray: the incoherent case
        oo, dd = oo[perm].contiguous(), dd[perm].contiguous()
        ss = torch.from_numpy(np.tile(seeds[sel], reps)[:a.n].astype(np.uint32).view(np.int32)).cuda()[perm].contiguous()
        out = torch.empty((a.n, 8), dtype=torch.float32, device="cuda")
        best = None
        for it in range(6):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            rc = F.lib().b200rt_intersect_batch_device(scene._h, C.c_void_p(oo.data_ptr()), C.c_void_p(dd.data_ptr()), C.c_void_p(ss.data_ptr()), a.n, 0.001, C.c_void_p(out.data_ptr()))
            assert rc == 0
            e1.record()
            e1.synchronize()
            ms = e0.elapsed_time(e1)
            if it and (best is None or ms < best):
                best = ms
        hits = int((out[:, 1].view(torch.int32) >= 0).sum().item())
        rows.append({"scene": name, "rays": a.n, "what": kind, "ms": best, "mrays_s": a.n / best / 1e3, "hit_fraction": hits / a.n,
                     "gbs_alg": a.n * 68 / best / 1e6})
        print(json.dumps(rows[-1]), flush=True)
    scene.close()
