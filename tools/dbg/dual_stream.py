"""Experiment: two wavefront pipelines (two contexts on two CUDA streams of ONE device, each with its own pool) render
the two halves of the sample range concurrently into the same accumulation buffer, so that one pipeline's launch
tails and shade passes are filled by the other's extend work.  usage: python tools/dbg/dual_stream.py [spp_total] [n_pipes]"""
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200.sharding import spp_range  # noqa: E402
from scenes_util import scene_builder  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 256
pipes = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sb = scene_builder("book2")
W = H = 800
accum = torch.zeros((H, W, 4), dtype=torch.float32, device="cuda:0")


def run(n_pipes, reps=4):
    ctxs, scenes, streams = [], [], []
    for k in range(n_pipes):
        c = rt.Context(0)
        s = torch.cuda.Stream(device=0)
        c.set_stream(s.cuda_stream)
        ctxs.append(c); streams.append(s); scenes.append(rt.Scene(c, sb))
    stats = [None] * n_pipes

    def work(k, begin):
        b, n = spp_range(k, n_pipes, spp, begin)
        p = rt.copy_params(sb.params, width=W, height=H, spp_begin=b, spp_count=n)
        stats[k] = scenes[k].render_device(sb.camera, p, accum.data_ptr())
    best = 1e9
    for rep in range(reps):
        accum.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        th = [threading.Thread(target=work, args=(k, rep * spp)) for k in range(n_pipes)]
        for t in th:
            t.start()
        for t in th:
            t.join()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if rep:
            best = min(best, dt)
    print("%d pipeline(s), %d spp total: %.2f ms wall -> %.1f Mpaths/s; per-pipeline device ms %s; mean %.5f" % (
        n_pipes, spp, best * 1e3, W * H * spp / best / 1e6, ["%.1f" % s.ms_total for s in stats], float(accum[..., :3].mean() / spp)), flush=True)
    for s in scenes:
        s.close()
    for c in ctxs:
        c.close()


run(1)
run(pipes)
run(1)
