"""Every BASELINE.json config as stated, on one GPU (C1 .. C5), each with

  * Mpaths/s and Mrays/s (CUDA events around capture_image; a warm-up call first),
  * the algorithmic FP32 and HBM fractions of SURVEY.md section 8(d) (operation counts from the counters build, tools/opcount.py),
  * the CPU baseline: the oracle's own construction of the same scene on all host cores at reduced spp (core count stated),
  * the north_star image gate: per-channel mean of the GPU image against the oracle's two-seed mean at >= 256 pixels a side.

  python tools/bench_configs.py [--spp 256] [--only C4] [--no-cpu] [--out profiles/r2_configs.json]
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import numpy as np  # noqa: E402
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder, oracle_scene, DATA  # noqa: E402
import opcount  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--spp", type=int, default=256, help="samples per pixel per timed call (capped by the config's own spp)")
ap.add_argument("--only", default="")
ap.add_argument("--no-cpu", action="store_true")
ap.add_argument("--no-image-gate", action="store_true")
ap.add_argument("--out", default="")
args = ap.parse_args()

# label, scene, width, height, config spp, integrator, tight oracle triangle boxes
CASES = [
    ("C1", "book-1 final (485 spheres), path", "book1", 400, 225, 100, "path", 0.0),
    ("C2", "Cornell book-3 (two rotated boxes), MixturePDF", "cornell_book3", 600, 600, 1000, "mixture", 0.0),
    ("C2n", "Cornell book-3, light-sample + shadow ray with MIS (light.rs:107-124 analogue)", "cornell_book3", 600, 600, 1000, "nee", 0.0),
    ("C3a", "data/scene_500.json (1,005 spheres, BVH)", "json:scene_500", 1920, 1080, 256, "path", 0.0),
    ("C3b", "data/scene_200_no_bvh.json (405 spheres, flat HitableList)", "json:scene_200_no_bvh", 1920, 1080, 256, "path", 0.0),
    ("C4", "book-2 final (1,409 primitives)", "book2", 800, 800, 10000, "path", 0.0),
    ("C5", "bun315.obj subdivided x3 (317,952 triangles) in the Cornell box, MixturePDF", "mesh", 1920, 1080, 4096, "mixture", 1e-6),
]
INTEG = {"path": F.INTEGRATOR_PATH, "mixture": F.INTEGRATOR_MIXTURE, "nee": F.INTEGRATOR_NEE}
peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))) if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else {"hbm_gbs": 6650.0}
ctx = rt.Context(0)
fp32_peak, _ = ctx.measure_fp32_peak()
threads = os.cpu_count() or 8
rows = []


def camera_for(name, sb, w, h):
    """The JSON files store aspect 1; SURVEY.md section 8(d) renders them at 16:9 — Camera::new again with the image's aspect."""
    if not name.startswith("json:"):
        return sb.camera
    c = json.load(open(os.path.join(DATA, name[5:] + ".json")))["camera"]
    v = lambda k: (c[k]["x"], c[k]["y"], c[k]["z"])  # noqa: E731
    return rt.Camera((v("look_from"), v("look_at")), v("vup"), c["vfov"], w / h, c["aperture"], c["focus_dist"]).c


for key, label, name, w, h, cfg_spp, integ, tri_pad in CASES:
    if args.only and key not in args.only.split(","):
        continue
    sb = scene_builder(name)
    cam = camera_for(name, sb, w, h)
    t0 = time.perf_counter()
    scene = rt.Scene(ctx, sb)
    t_create = time.perf_counter() - t0
    n = min(args.spp, cfg_spp)
    if key == "C3b":
        n = max(1, n // 8)  # the brute-force arm is ~13x slower per path
    p = rt.copy_params(sb.params, width=w, height=h, spp_count=n, integrator=INTEG[integ])
    scene.render(cam, p)  # warm-up (allocates the queue pool)
    best = None
    for rep in range(2):
        acc, st = scene.render(cam, rt.copy_params(p, spp_begin=(rep + 1) * n))
        if best is None or st.ms_total < best.ms_total:
            best = st
    st = best
    rays = st.extend_rays + st.shadow_rays
    sec = st.ms_total * 1e-3
    info = scene.info()
    row = {"config": key, "what": label, "resolution": "%dx%d" % (w, h), "config_spp": cfg_spp, "spp_per_call": n, "integrator": integ,
           "mpaths_s": st.paths / sec / 1e6, "mrays_s": rays / sec / 1e6, "rays_per_path": rays / st.paths, "ms_per_call": st.ms_total,
           "primitives": info["device_prims"], "bvh_nodes": info["bvh_nodes"], "scene_create_ms": t_create * 1e3, "pipelines": st.pipelines,
           "hbm_frac_alg": (128.0 * st.extend_rays + 96.0 * st.shadow_rays + 32.0 * st.paths) / sec / 1e9 / peaks["hbm_gbs"]}
    fpp, detail = opcount.algorithmic_flops(name, w, h, integ, spp=4 if key != "C3b" else 1)
    if fpp:
        row.update(flops_per_path=fpp, tflops_alg=fpp * st.paths / sec / 1e12, fp32_frac_alg=fpp * st.paths / sec / 1e12 / fp32_peak, op_counts=detail["counts"])
    else:
        row["fp32_note"] = detail
    if not args.no_cpu:
        orc = oracle_scene(name, tri_pad=tri_pad)
        import oracle_ffi
        cspp = 1
        pc = oracle_ffi.oracle_py.copy_params(p, width=w, height=h, spp_count=cspp, seed=7)
        _, ext, shd, csec = orc.render(cam, pc, n_jobs=max(64, 4 * threads), n_threads=threads)
        if csec < 3.0 and key != "C3b":  # scale the sample to a few seconds
            cspp = int(min(16, max(1, 6.0 / max(csec, 1e-3))))
            pc = oracle_ffi.oracle_py.copy_params(p, width=w, height=h, spp_count=cspp, seed=8)
            _, ext, shd, csec = orc.render(cam, pc, n_jobs=max(64, 4 * threads), n_threads=threads)
        row.update(cpu_mpaths_s=w * h * cspp / csec / 1e6, cpu_mrays_s=(ext + shd) / csec / 1e6, cpu_cores=threads,
                   cpu_sample="%dx%d x %d spp, %.1f s%s" % (w, h, cspp, csec, ", oracle triangle boxes padded by 1e-6 instead of mesh.rs:33-42's 0.1 (same answers, usable BVH)" if tri_pad else ""))
        if not args.no_image_gate:
            # image gate at >= 256 pixels a side, same aspect: GPU at 4 x the oracle's samples against the oracle's two-seed mean
            gh = 256 if h <= w else int(round(256 * h / w))
            gw = int(round(gh * w / h)) if h <= w else 256
            gspp = {"C3b": 8, "C5": 16}.get(key, 32)
            pg = rt.copy_params(p, width=gw, height=gh, spp_count=gspp * 8, spp_begin=0)
            camg = camera_for(name, sb, gw, gh) if name.startswith("json:") else cam
            g, _ = scene.render(camg, pg)
            gpu_mean = (g[..., :3].astype(np.float64) / (gspp * 8))
            pa = oracle_ffi.oracle_py.copy_params(pg, spp_count=gspp, seed=101)
            pb = oracle_ffi.oracle_py.copy_params(pg, spp_count=gspp, seed=202)
            a, _, _, _ = orc.render(camg, pa, n_jobs=max(64, 4 * threads), n_threads=threads)
            b, _, _, _ = orc.render(camg, pb, n_jobs=max(64, 4 * threads), n_threads=threads)
            a, b = a / gspp, b / gspp
            ref = 0.5 * (a + b)
            cap = np.percentile(ref, 99.9) * 4 + 1e-3  # clamp fireflies the same way on both sides
            gm, rm = np.minimum(gpu_mean, cap).mean(axis=(0, 1)), np.minimum(ref, cap).mean(axis=(0, 1))
            sigma = np.sqrt(((np.minimum(a, cap) - np.minimum(b, cap)) ** 2).mean(axis=(0, 1)) / 2.0 / (gw * gh)) / np.sqrt(2.0)
            row.update(image_gate={"resolution": "%dx%d" % (gw, gh), "gpu_spp": gspp * 8, "oracle_spp": "2 x %d" % gspp,
                                   "mean_err_pct": (100 * np.abs(gm - rm) / np.maximum(rm, 1e-9)).round(3).tolist(),
                                   "oracle_noise_pct_1sigma": (100 * sigma / np.maximum(rm, 1e-9)).round(3).tolist()})
    rows.append(row)
    print(json.dumps(row), flush=True)
    scene.close()

if args.out:
    with open(args.out, "w") as f:
        json.dump({"fp32_peak_tflops_measured": fp32_peak, "hbm_gbs": peaks["hbm_gbs"], "rows": rows}, f, indent=1)
