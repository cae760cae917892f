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
FULL_JOB_SPP = 10000        # BASELINE.json: "book-2 final scene 800x800, 10k spp at 1/2/4/8 B200" — the default step IS that job,
                            # split over the ranks by sample index (strong scaling: the total work is fixed as N grows)
SPP_PER_STEP = 512          # --scaling weak: per GPU and step (one capture_image call).  A call drains its pool once at the end, so
                            # throughput rises with its size: 128 / 256 / 512 / 1250 spp per call = 1477 / 1515 / 1529 / 1535 Mpaths/s
METRIC = "Mpaths/s (device-timed) on book-2 final scene"
UNIT = "Mpaths/s"
# extend kernel, per ray: the 32-byte ray stream read (o, time | d, medium seed) + the 16-byte hit written (t as f64, prim, rank)
EXTEND_BYTES_PER_RAY = 48
SEGMENT_BYTES_PIPELINE = 128  # ray 32 + hit 16 + state 16, written once and read once
PATH_BYTES = 32               # one float4 framebuffer read-modify-write per path
# ncu --set full of the dominant kernel (constants of one committed capture; bench.py measures everything else live)
NCU = json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json")))


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
    """The oracle (C++ f64 restatement of the reference) on the host cores: returns (Mpaths/s, Mrays/s, seconds).
    The scene is the oracle's OWN construction of the book-2 fixture (oracle/oracle_scenes.inc: the reference's object graph
    with nested BVHNode::new, Transform over the sphere BVH, ConstantMedium over its boundary); nothing of the product is
    imported or mapped on this path (oracle/oracle_py.py is standalone)."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle_py
    orc = oracle_py.OracleScene.make("book2", seed=1, earth=earth_image())
    p = oracle_py.copy_params(orc.params, width=WIDTH, height=HEIGHT, spp_count=spp, seed=seed)
    _, ext, shd, sec = orc.render(orc.camera, p, n_jobs=max(64, 4 * threads), n_threads=threads)
    paths = WIDTH * HEIGHT * spp
    return paths / sec / 1e6, (ext + shd) / sec / 1e6, sec


def algorithmic_flops():
    """flops_alg = sum(counts x C), the counts measured on this box by the -DB200RT_COUNTERS twin of the
    library on the benchmark scene (tools/opcount.py, a child process; never on the timed path)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import opcount
    return opcount.algorithmic_flops("book2", WIDTH, HEIGHT, "path")


def run_sppm(args):
    """--workload sppm: what the reference's main() renders (main.rs:33-53): scene.rs:16-112 at 800x800, SPPMIntegrator::new =
    50 iterations x 500,000 photons (photon_mapper.rs:148-149), then capture_image at 256 spp over SPPMIntegrator::sample_ray.
    One step = one such frame through b200rt_render_sppm (host buffers in and out: the timed region is the public call).
    cpu_baseline: the oracle restatement on bounded samples (one iteration at two small image sizes + a few final-pass samples)
    scaled to the full job, which is exact for this algorithm (every iteration, pixel and sample costs the same on average);
    its k-nearest is an exhaustive scan standing in for kd-tree 0.4.1, so only its photon pass and final pass are like for like."""
    import time as _t
    import numpy as np
    import rust_raytracer_b200 as rt
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from scenes_util import scene_builder, oracle_scene
    W = H = 800
    ITER, PHOTONS, SPP = 50, 500000, 256
    sb = scene_builder("cornell_ref")
    ctx = rt.Context(0)
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=W, height=H, spp_count=SPP, seed=1)
    for _ in range(max(args.warmup, 1)):
        scene.render_sppm(sb.camera, p, iterations=ITER, photons_per_iter=PHOTONS)
    sampler = ClockSampler(0)
    sampler.start()
    ms, wall, last = [], [], None
    for i in range(args.steps):
        t0 = _t.perf_counter()
        acc, st = scene.render_sppm(sb.camera, rt.copy_params(p, seed=2 + i), iterations=ITER, photons_per_iter=PHOTONS)
        wall.append(_t.perf_counter() - t0)
        ms.append(st.ms_total)
        last = st
    _t.sleep(0.3)
    sampler.stop()
    st = last
    cpu = None
    if not args.no_cpu_baseline:
        threads = os.cpu_count() or 8
        orc = oracle_scene("cornell_ref")
        # bounded sample: ONE iteration with the full 500,000 photons on an 8x8 image (the photon pass; the 64 eye look-ups of the
        # exhaustive-scan checker are ~0.05 s inside it), then the same with 8192 more final-pass samples per pixel

        def cpu_run(size, spp):
            t0 = _t.perf_counter()
            orc.sppm_render(sb.camera, rt.copy_params(p, width=size, height=size, spp_count=spp), iterations=1, photons_per_iter=PHOTONS, n_threads=threads)
            return _t.perf_counter() - t0
        cpu_run(8, 1)  # warm-up (page-in, thread pool)
        t_a, t_c = cpu_run(8, 1), cpu_run(8, 8193)
        per_photon = t_a / PHOTONS
        per_sample = max(t_c - t_a, 0.0) / (8192 * 64)
        est_like_for_like = ITER * PHOTONS * per_photon + SPP * W * H * per_sample
        cpu = {"value": est_like_for_like * 1e3, "unit": "ms per frame (photon passes + final pass only)", "cores": threads, "kind": "port",
               "sample": "1 iteration x %d photons at 8x8 with 1 spp (%.2f s) and with 8193 spp (%.2f s), scaled to 50 x 500,000 photons and 800x800x256 samples" % (PHOTONS, t_a, t_c),
               "us_per_photon": per_photon * 1e6, "us_per_final_sample": per_sample * 1e6,
               "note": "the oracle's eye pass replaces kd-tree 0.4.1 by an exhaustive scan (a checker, not a baseline) and is left out of this figure; the GPU time beside it is ms_photons + ms_final"}
    n_emit, n_eye, n_final = ITER * PHOTONS, ITER * W * H, SPP * W * H
    out = {"metric": "SPPM frame of the reference's main() (cornell_ref 800x800, 50 x 500,000 photons, 256 spp)", "value": float(np.mean(ms)), "unit": "ms per frame",
           "n_gpus": 1, "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": float(np.mean(ms)), "higher_is_better": False, "scaling": "n/a (one device)", "vs_baseline": None,
           "dtype": "f32 (f64 hit confirmation)", "data": "synthetic",
           "config": {"workload": "SPPMIntegrator::new + capture_image (photon_mapper.rs:139-236, 327-365; main.rs:33-53)", "iterations": ITER, "photons_per_iteration": PHOTONS, "spp": SPP},
           "stage_ms": {"photons": st.ms_photons, "grid": st.ms_grid, "eye": st.ms_eye, "final": st.ms_final, "total": st.ms_total,
                        "note": "up to 7 photon passes run ahead of the eye pass on their own streams: photons is the sum of the passes' own event pairs (they overlap each other and the eye passes), grid / eye / final are on the context stream"},
           "rates": {"Mphotons_emitted_per_s": n_emit / st.ms_photons / 1e3, "Mphoton_rays_per_s": st.photon_rays / st.ms_photons / 1e3,
                     "Meye_updates_per_s": n_eye / st.ms_eye / 1e3, "Mfinal_paths_per_s": n_final / st.ms_final / 1e3},
           "photons_stored": {"global": int(st.global_photons), "caustic": int(st.caustic_photons)},
           "roofline": {"bound": "fp32_issue", "kernel": "k_sppm_gather (SPPM::update radius gather: 49 of the 50 eye passes)", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None,
                        "issue": {"issue_active_pct": 50.0, "lanes_per_instruction": 18.5, "fma_pipe_pct": 23.2, "l1_hit_pct": 76.9, "l2_hit_pct": 85.6, "dram_pct_of_peak": 0.9,
                                  "source": "profiles/r2_sppm_kernels_final.txt (ncu --set full, iteration 1 of tools/bench_sppm.py)"},
                        "note": "SURVEY.md section 8(d) defines no algorithmic flop count for the photon-map look-ups, so no fraction is claimed; the kernel is issue / latency bound (the maps are L2-resident: 0.9 % of HBM), and the frame is bounded by the eye passes (stage_ms.eye) with the photon passes hidden under them (7 in flight)"},
           "cpu_baseline": cpu,
           "e2e": {"value": float(np.mean(wall)) * 1e3, "unit": "ms per frame", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": W * H * 16,
                   "what": "host clock around b200rt_render_sppm (image returned to a host buffer)"},
           "gpu_launches": None, "clocks": sampler.summary(), "device": ctx.device_info()}
    print(json.dumps(out))
    return 0


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
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default): a step is the full 10,000-spp job split over the ranks; weak: --spp samples per GPU per step")
    ap.add_argument("--spp", type=int, default=0, help="strong: total spp of the job (default 10000); weak: spp per GPU per step (default 512)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--workload", default="path", choices=["path", "sppm"], help="path: the BASELINE.json metric (default); sppm: the reference's own main()")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    if args.workload == "sppm":
        return run_sppm(args)
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
    json_out = sys.stdout
    if world > 1:
        # NCCL's version / debug lines must not precede the JSON line on stdout (it prints "NCCL version ..." to fd 1 from native
        # code when the first communicator comes up): fd 1 points at stderr for the whole run, the JSON line goes to the saved fd
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        sys.stdout.flush()
        json_out = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = torch.device("cuda", local_rank)
    strong = args.scaling == "strong"
    job_spp = args.spp or (FULL_JOB_SPP if strong else SPP_PER_STEP)
    if strong:  # contiguous sample ranges of the fixed job, remainder to the first ranks
        spp = job_spp // world + (1 if rank < job_spp % world else 0)
        rank_first = rank * (job_spp // world) + min(rank, job_spp % world)
        step_spp = job_spp
    else:
        spp, rank_first, step_spp = job_spp, rank * job_spp, job_spp * world

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
        begin = step_index * step_spp + rank_first
        p = rt.copy_params(params, spp_begin=begin)
        accum.zero_()
        st = scene.render_device(sb.camera, p, accum.data_ptr())
        reduce_accum(accum, dst=0)
        for k in ("paths", "extend_rays", "shadow_rays", "kernel_launches", "wavefront_iterations"):
            stats_acc[k] = stats_acc.get(k, 0) + getattr(st, k)
        for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow", "ms_total", "ms_generate_sum", "ms_extend_sum", "ms_shade_sum", "ms_shadow_sum"):
            stats_acc[k] = stats_acc.get(k, 0.0) + getattr(st, k)
        stats_acc["pipelines"] = st.pipelines

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
    e2e_steps = 3
    e2e_marks = []
    sync_all()
    t0 = time.perf_counter()
    for i in range(-1, e2e_steps):
        if i == 0:  # iteration -1 is an untimed warm-up: the GPU idled (and clocked down) while the stage statistics were gathered
            sync_all()
            t0 = time.perf_counter()
        sc2 = rt.Scene(ctx, sb)  # World::new: validate, bake, build the BVH, upload (H2D)
        begin = (args.warmup + args.steps + 1 + i) * step_spp + rank_first
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
    e2e_value = WIDTH * HEIGHT * step_spp * e2e_steps / float(e2e_t.item()) / 1e6

    # ---- one step with ONE pipeline (rank 0): clean per-kernel durations (with two concurrent pipelines the stage intervals overlap)
    single = None
    if rank == 0 and args.pipeline == "wavefront":
        os.environ["B200RT_WF_PIPES"] = "1"
        try:
            single = {}
            torch.cuda.synchronize()
            p1 = rt.copy_params(params, spp_begin=(args.warmup + args.steps + e2e_steps + 2) * step_spp, spp_count=min(spp, 512))
            accum.zero_()
            st1 = scene.render_device(sb.camera, p1, accum.data_ptr())
            for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow", "ms_total"):
                single[k] = getattr(st1, k)
            single.update(spp=min(spp, 512), iterations=st1.wavefront_iterations, extend_rays=int(st1.extend_rays), paths=int(st1.paths))
        finally:
            os.environ.pop("B200RT_WF_PIPES", None)

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        fp32_tflops, eff_mhz = ctx.measure_fp32_peak()
        flops_per_path, flop_detail = algorithmic_flops() if args.pipeline == "wavefront" else (None, "megakernel")
        n_iter = max(1, stats["wavefront_iterations"])
        rays_per_path = (ext + shd) / paths
        # the dominant kernel, k_wf_extend, from the single-pipeline step: its launches do not overlap anything there
        if single and single["iterations"] > 0 and single["ms_extend"] > 0:
            ext_launch_ms = single["ms_extend"] / single["iterations"]
            ext_rays_launch = single["extend_rays"] / single["iterations"]
            ext_share = single["ms_extend"] / single["ms_total"]
            paths_launch = single["paths"] / single["iterations"]
        else:
            ext_launch_ms = ext_rays_launch = ext_share = paths_launch = None
        ext_flops_path = (flop_detail["flop_per_path"]["box_tests"] + flop_detail["flop_per_path"]["primitive_tests"]) if flops_per_path else None
        ext_tflops = ext_flops_path * paths_launch / (ext_launch_ms * 1e-3) / 1e12 if (ext_flops_path and ext_launch_ms) else None
        ext_gbs = EXTEND_BYTES_PER_RAY * ext_rays_launch / (ext_launch_ms * 1e-3) / 1e9 if ext_launch_ms else None
        lanes = NCU["extend"]["lanes_per_instruction"]
        pipeline_bytes = SEGMENT_BYTES_PIPELINE * ext + 96 * shd + PATH_BYTES * paths
        roofline = {
            "bound": "fp32_issue", "kernel": "k_wf_extend" if args.pipeline == "wavefront" else "k_megakernel",
            "achieved": ext_tflops, "peak": fp32_tflops, "unit": "TFLOP/s", "frac": (ext_tflops / fp32_tflops) if ext_tflops else None,
            "fp32_frac": (ext_tflops / fp32_tflops) if ext_tflops else None,
            "peak_source": "measured live: k_fma_peak (8 independent FMA chains per thread, 8 x 256 threads per SM), %.0f MHz effective" % eff_mhz,
            "what": "SURVEY.md section 8(d): algorithmic flops of the launch = (25 flop x ray-box slab tests + per-type constants x primitive tests), the operation counts measured on this box by the -DB200RT_COUNTERS twin of the library on the same scene, x paths per launch / the kernel's average launch duration in a single-pipeline step (CUDA events)",
            "flops_per_path_kernel": ext_flops_path, "paths_per_launch": paths_launch, "rays_per_launch": ext_rays_launch, "avg_launch_ms": ext_launch_ms,
            "kernel_share_of_step": ext_share,
            "issue": {"issue_active_pct": NCU["extend"]["issue_active_pct"], "fma_pipe_pct": NCU["extend"]["fma_pipe_pct"], "lanes_per_instruction": lanes,
                      "lane_weighted_fma_pipe_pct": NCU["extend"]["fma_pipe_pct"] * lanes / 32.0, "alu_pipe_pct": NCU["extend"]["alu_pipe_pct"],
                      "fp64_pipe_pct": NCU["extend"]["fp64_pipe_pct"], "l1_hit_pct": NCU["extend"]["l1_hit_pct"], "l2_hit_pct": NCU["extend"]["l2_hit_pct"],
                      "source": NCU["source"]},
            "traffic": NCU["extend"]["dram_bytes_per_ray"] * ext_rays_launch if ext_rays_launch else None,
            "traffic_note": "ncu dram__bytes_read + dram__bytes_write per ray of one launch (%.1f B against %d algorithmic) x rays of the average launch" % (NCU["extend"]["dram_bytes_per_ray"], EXTEND_BYTES_PER_RAY),
            "hbm": {"achieved": ext_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": (ext_gbs / peaks["hbm_gbs"]) if ext_gbs else None, "bytes_per_ray": EXTEND_BYTES_PER_RAY,
                    "peak_source": "MEASURED_PEAKS.json (%s)" % peak_kind, "note": "secondary roof: the scene is L2-resident, extend streams 32 B in and 16 B out per ray"},
            "pipeline": {"fp32_frac": (flops_per_path * paths / total_s / 1e12 / fp32_tflops) if flops_per_path else None,
                         "tflops_algorithmic": (flops_per_path * paths / total_s / 1e12) if flops_per_path else None,
                         "flops_per_path": flops_per_path,
                         "hbm_frac": pipeline_bytes / total_s / 1e9 / peaks["hbm_gbs"], "hbm_gbs": pipeline_bytes / total_s / 1e9,
                         "bytes": "128 B per extend ray + 96 B per shadow ray + 32 B per path (SURVEY.md section 8(d))"},
            "operation_counts": flop_detail,
        }
        cpu = None
        if not args.no_cpu_baseline and world == 1:
            threads = os.cpu_count() or 8
            _, _, sec = cpu_baseline_run(1, threads)
            cspp = max(1, min(32, int(12.0 / max(sec, 1e-3))))
            mp, mr, sec = cpu_baseline_run(cspp, threads)
            cpu = {"value": mp, "unit": UNIT, "cores": threads, "kind": "port", "mrays_per_s": mr,
                   "sample": "800x800 x %d spp (%.1f s), oracle/ C++ f64 restatement on its own object graph (oracle/oracle_scenes.inc), %d threads" % (cspp, sec, threads)}
        n_pipes = int(stats.get("pipelines", 1))
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_s * 1e3 / args.steps, "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None,
            "dtype": "f32 (f64 hit confirmation)", "data": "synthetic",
            "config": {"workload": "book-2 final scene 800x800 (BASELINE.json configs[3]): 1409 primitives incl. 400 boxes, 1000 instanced spheres, moving sphere, 2 constant media, Perlin + earthmap textures",
                       "step": ("the full %d-spp job, split over the %d rank(s) by sample index (%d spp on this rank)" % (job_spp, world, spp)) if strong else ("%d spp per GPU" % spp),
                       "spp_per_step_total": step_spp, "spp_per_step_per_gpu": spp, "full_job_spp": FULL_JOB_SPP, "max_depth": 50, "integrator": "path", "pipeline": args.pipeline,
                       "sharding": "spp ranges per rank + NCCL sum-reduce of the accumulation buffer per step",
                       "concurrent_pipelines": n_pipes,
                       "l2": "256 MB write between timed steps; the wavefront queues (2^28 positions x 144 B of ray/state/hit streams = 38.7 GB) exceed L2"},
            "mrays_per_s": mrays, "rays_per_path": rays_per_path,
            "stage_ms_per_step": {k: stats[k] / args.steps for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow", "ms_total")},
            "stage_ms_sum_per_step": {k: stats[k + "_sum"] / args.steps for k in ("ms_generate", "ms_extend", "ms_shade", "ms_shadow")},
            "stage_ms_single_pipeline": single,
            "stage_ms_note": "capture_image runs %d concurrent wavefront pipeline(s) (own stream each). stage_ms_per_step: length of the UNION of a stage's CUDA-event intervals over the pipelines (they overlap: the sum exceeds ms_total). stage_ms_sum_per_step: every launch's own event-pair duration added up (a launch queued behind the other pipeline's blocks counts its wait). stage_ms_single_pipeline: one extra, untimed step with ONE pipeline, where stages neither overlap nor queue: the per-kernel durations the roofline uses" % n_pipes,
            "wavefront_iterations_per_step": n_iter / args.steps,
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
        json_out.write(json.dumps(out) + "\n")
        json_out.flush()
    if world > 1:
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
