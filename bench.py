#!/usr/bin/env python
"""bench.py — Mpaths/s on the book-2 final scene (BASELINE.json metric), 1..8 B200.

A step = one pass of the hot path (capture_image) over 800x800 pixels x SPP_PER_STEP samples per GPU.
Each rank renders a disjoint global sample range (spp sharding) and one NCCL sum-reduce of the
float accumulation buffer closes the step (inside the timed region).  `value` = paths of all ranks / device time.

  python bench.py --gpus 1 --steps 5 --warmup 3
  python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...      # the CPU restatement of the reference (oracle/), all host threads
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WIDTH, HEIGHT = 800, 800
SPP_PER_STEP = 512          # per GPU and step = one capture_image call; the full job is 10,000 spp.  A call drains its pool once at the
                            # end, so throughput rises with its size: 128 / 256 / 512 / 1250 spp per call = 1477 / 1515 / 1529 / 1535 Mpaths/s
FULL_JOB_SPP = 10000
METRIC = "Mpaths/s (device-timed) on book-2 final scene"
UNIT = "Mpaths/s"
# extend kernel, per ray: the 32-byte ray stream read (o, time | d, medium seed) + the 16-byte hit written (t as f64, prim, rank)
EXTEND_BYTES_PER_RAY = 48
SEGMENT_BYTES_PIPELINE = 128  # ray 32 + hit 16 + state 16, written once and read once
# ncu --set full, k_wf_extend, launch of 18,292,061 rays: dram__bytes_read.sum 0.5996 GB + dram__bytes_write.sum 0.3108 GB
# (profiles/r1_final2_extend_shade_full.txt) -> DRAM bytes per ray; issue/lane figures from the same capture
EXTEND_DRAM_BYTES_PER_RAY_NCU = (0.599593e9 + 310.840576e6) / 18292061
EXTEND_NCU = {"issue_active_pct": 76.4, "alu_pipe_pct": 48.6, "fma_pipe_pct": 27.3, "fp64_pipe_pct": 6.3, "lanes_per_instruction": 14.4,
              "warp_instructions_per_32_rays": 3144, "l1_hit_pct": 86.3, "l2_hit_pct": 73.6, "dram_pct_of_peak": 5.4,
              "shade_issue_active_pct": 50.8, "shade_dram_pct_of_peak": 36.0, "shade_dram_bytes_per_ray": 125.2,
              "source": "profiles/r1_final2_extend_shade_full.txt"}


def earth_image():
    import numpy as np
    try:
        from PIL import Image
        return np.asarray(Image.open(os.path.join(ROOT, "data", "earthmap.jpg")).convert("RGB"), dtype=np.uint8)
    except Exception:
        return np.load(os.path.join(ROOT, "tests", "golden", "earthmap_256x128.npy"))


def build_scene_builder():
    import rust_raytracer_b200 as rt
    sb = rt.SceneBuilder().make_scene("book2", seed=1, earth=earth_image())
    params = rt.copy_params(sb.params, width=WIDTH, height=HEIGHT, spp_count=SPP_PER_STEP)
    return sb, params


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt, self.proc = index, [], threading.Event(), None

    def run(self):
        q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q, "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                if self._stop_evt.is_set():
                    break
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        self._stop_evt.set()
        if self.proc:
            try:
                self.proc.terminate()
            except Exception:
                pass

    def summary(self):
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx or None, "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0}, "fallback"


def cpu_baseline_run(spp, threads, seed=1):
    """The oracle (C++ f64 restatement of the reference) on the host cores: returns (Mpaths/s, Mrays/s, seconds)."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import rust_raytracer_b200 as rt
    from oracle_ffi import OracleScene
    sb, params = build_scene_builder()
    orc = OracleScene(sb)
    p = rt.copy_params(params, spp_count=spp, seed=seed)
    _, ext, shd, sec = orc.render(sb.camera, p, n_jobs=max(64, 4 * threads), n_threads=threads)
    paths = WIDTH * HEIGHT * spp
    return paths / sec / 1e6, (ext + shd) / sec / 1e6, sec


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    threads = os.cpu_count() or 8
    # bounded sample per step: ~2-4 s of CPU work, scaled after a probe
    mp, _, sec = cpu_baseline_run(1, threads)
    spp = max(1, min(16, int(3.0 / max(sec, 1e-3))))
    for _ in range(args.warmup):
        cpu_baseline_run(spp, threads)
    rays, dt = 0.0, 0.0
    for k in range(args.steps):
        _, mr, s = cpu_baseline_run(spp, threads, seed=100 + k)
        rays += mr * s
        dt += s  # capture_image only (steady_clock inside the oracle), scene construction excluded as on the GPU side
    value = WIDTH * HEIGHT * spp * args.steps / dt / 1e6
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "book-2 final scene 800x800 (BASELINE.json configs[3]), CPU restatement of the reference, %d spp per step" % spp,
                   "spp_per_step": spp, "max_depth": 50, "integrator": "path"},
        "mrays_per_s": rays / dt,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": "800x800 x %d spp per step, %d steps, %d threads, 64+ row-band jobs (camera.rs:74-111 decomposition)" % (spp, args.steps, threads)},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200rt")
    ap.add_argument("--pipeline", default="wavefront", choices=["wavefront", "megakernel"])
    ap.add_argument("--spp", type=int, default=SPP_PER_STEP)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    args.warmup = max(args.warmup, 3)

    import numpy as np
    import torch
    import torch.distributed as dist
    import rust_raytracer_b200 as rt
    from rust_raytracer_b200 import _ffi as F
    from rust_raytracer_b200.sharding import reduce_accum

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product path has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")  # NCCL's version / debug lines must not precede the JSON line on stdout
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    spp = args.spp

    ctx = rt.Context(local_rank)
    stream = torch.cuda.current_stream()
    ctx.set_stream(stream.cuda_stream)
    sb, params = build_scene_builder()
    params = rt.copy_params(params, spp_count=spp, pipeline=F.PIPELINE_MEGAKERNEL if args.pipeline == "megakernel" else F.PIPELINE_WAVEFRONT)
    scene = rt.Scene(ctx, sb)
    info = scene.info()
    accum = torch.zeros((HEIGHT, WIDTH, 4), dtype=torch.float32, device=dev)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def one_step(step_index, stats_acc):
        # global sample range of this rank for this step: disjoint across ranks and steps
        begin = (step_index * world + rank) * spp
        p = rt.copy_params(params, spp_begin=begin)
        accum.zero_()
        st = scene.render_device(sb.camera, p, accum.data_ptr())
        reduce_accum(accum, dst=0)
        for k in ("paths", "extend_rays", "shadow_rays", "kernel_launches", "wavefront_iterations"):
            stats_acc[k] = stats_acc.get(k, 0) + getattr(st, k)
        for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow", "ms_total"):
            stats_acc[k] = stats_acc.get(k, 0.0) + getattr(st, k)

    warm = {}
    for i in range(args.warmup):
        one_step(i, warm)
    sync_all()

    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    stats = {}
    step_ms = []
    sync_all()
    for i in range(args.steps):
        flush.fill_(i & 0xFF)  # L2 flush between timed iterations (outside the event pair)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        one_step(args.warmup + i, stats)
        e1.record(stream)
        e1.synchronize()
        step_ms.append(e0.elapsed_time(e1))
    sync_all()
    if rank == 0:
        time.sleep(0.3)
        sampler.stop()
    total_ms = torch.tensor([sum(step_ms)], dtype=torch.float64, device=dev)
    counts = torch.tensor([stats["paths"], stats["extend_rays"], stats["shadow_rays"]], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(total_ms, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    total_s = float(total_ms.item()) * 1e-3
    paths, ext, shd = (float(x) for x in counts.tolist())
    value = paths / total_s / 1e6
    mrays = (ext + shd) / total_s / 1e6

    # ---- end to end through the public API with HOST buffers: scene tables H2D + render + reduce + image D2H
    host_img = torch.empty((HEIGHT, WIDTH, 4), dtype=torch.float32).pin_memory()
    e2e_steps = max(3, min(args.steps, 5))
    e2e_marks = []
    sync_all()
    t0 = time.perf_counter()
    for i in range(-1, e2e_steps):
        if i == 0:  # iteration -1 is an untimed warm-up: the GPU idled (and clocked down) while the stage statistics were gathered
            sync_all()
            t0 = time.perf_counter()
        sc2 = rt.Scene(ctx, sb)  # World::new: validate, bake, build the BVH, upload (H2D)
        begin = ((args.warmup + args.steps + i) * world + rank) * spp
        accum.zero_()
        sc2.render_device(sb.camera, rt.copy_params(params, spp_begin=begin), accum.data_ptr())
        reduce_accum(accum, dst=0)
        if rank == 0:
            host_img.copy_(accum, non_blocking=True)
        torch.cuda.synchronize()
        sc2.close()
        e2e_marks.append(time.perf_counter())
    sync_all()
    e2e_s = time.perf_counter() - t0
    e2e_t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(e2e_t, op=dist.ReduceOp.MAX)
    e2e_value = WIDTH * HEIGHT * spp * world * e2e_steps / float(e2e_t.item()) / 1e6

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        n_ext_launches = max(1, stats["wavefront_iterations"]) if args.pipeline == "wavefront" else args.steps
        ext_ms = stats["ms_extend"]
        rays_per_launch = stats["extend_rays"] / n_ext_launches
        achieved = EXTEND_BYTES_PER_RAY * stats["extend_rays"] / (ext_ms * 1e-3) / 1e9 if ext_ms > 0 else None
        fp32_tflops, eff_mhz = ctx.measure_fp32_peak()
        roofline = {
            "bound": "hbm", "kernel": "k_wf_extend" if args.pipeline == "wavefront" else "k_megakernel",
            "achieved": achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": (achieved / peaks["hbm_gbs"]) if achieved else None,
            "peak_source": "MEASURED_PEAKS.json (%s)" % peak_kind,
            "traffic": EXTEND_DRAM_BYTES_PER_RAY_NCU * rays_per_launch if args.pipeline == "wavefront" else None,
            "traffic_note": "ncu dram bytes per ray of one launch (%.1f B against %d algorithmic: the ray stream is read whole, the rest is local-memory stack traffic) x rays of the average launch" % (EXTEND_DRAM_BYTES_PER_RAY_NCU, EXTEND_BYTES_PER_RAY),
            "bytes_per_ray": EXTEND_BYTES_PER_RAY, "rays_per_launch": rays_per_launch, "launches": n_ext_launches,
            "avg_launch_ms": ext_ms / n_ext_launches, "kernel_share_of_step": ext_ms / max(stats["ms_total"], 1e-9),
            "note": "BVH traversal is instruction-issue bound under divergence, not HBM bound (scene is L2-resident): see issue, fp32 and DESIGN.md; the duration is the union of the extend launches' event intervals over the two concurrent pipelines (time during which an extend launch was in flight)",
            "issue": EXTEND_NCU,
            "pipeline_bytes_per_segment": SEGMENT_BYTES_PIPELINE,
            "pipeline_hbm_frac": SEGMENT_BYTES_PIPELINE * (ext + shd) / total_s / 1e9 / peaks["hbm_gbs"],
        }
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 8
            _, _, sec = cpu_baseline_run(1, threads)
            cspp = max(1, min(32, int(12.0 / max(sec, 1e-3))))
            mp, mr, sec = cpu_baseline_run(cspp, threads)
            cpu = {"value": mp, "unit": UNIT, "cores": threads, "kind": "port", "mrays_per_s": mr,
                   "sample": "800x800 x %d spp (%.1f s), oracle/ C++ f64 restatement, %d threads" % (cspp, sec, threads)}
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_s * 1e3 / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (f64 hit confirmation)", "data": "synthetic",
            "config": {"workload": "book-2 final scene 800x800 (BASELINE.json configs[3]): 1409 primitives incl. 400 boxes, 1000 instanced spheres, moving sphere, 2 constant media, Perlin + earthmap textures",
                       "spp_per_step_per_gpu": spp, "full_job_spp": FULL_JOB_SPP, "max_depth": 50, "integrator": "path", "pipeline": args.pipeline,
                       "sharding": "spp ranges per rank + NCCL sum-reduce of the accumulation buffer per step",
                       "concurrent_pipelines": int(os.environ.get("B200RT_WF_PIPES", "2")),
                       "l2": "256 MB write between timed steps; the wavefront queues (2^26 positions x 144 B of ray/state/hit streams = 9.7 GB) exceed L2"},
            "mrays_per_s": mrays, "rays_per_path": (ext + shd) / paths,
            "stage_ms_per_step": {k: stats[k] / args.steps for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow", "ms_total")},
            "stage_ms_note": "capture_image runs 2 concurrent wavefront pipelines (half of the samples each, own stream): a stage time is the length of the union of that stage's CUDA-event intervals over the pipelines, queueing behind the other pipeline's kernels included, so the stages overlap and add up to more than ms_total",
            "wavefront_iterations_per_step": stats["wavefront_iterations"] / args.steps,
            "roofline": roofline,
            "fp32": {"peak_tflops_measured": fp32_tflops, "effective_sm_mhz": eff_mhz, "how": "k_fma_peak: 8 independent FMA chains/thread, 8 blocks x 256 threads per SM"},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": info["h2d_bytes"], "d2h_bytes_per_step": WIDTH * HEIGHT * 16, "steps": e2e_steps,
                    "ms_per_step_list": [round((b - a) * 1e3, 2) for a, b in zip(e2e_marks[:-1], e2e_marks[1:])],
                    "what": "World::new (validate + bake + BVH build + upload) + capture_image + reduce + image D2H to pinned host"},
            "gpu_launches": int(stats["kernel_launches"]),
            "clocks": sampler.summary(),
            "device": ctx.device_info(),
            "scene": info,
        }
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
