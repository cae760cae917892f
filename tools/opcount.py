"""Algorithmic flops of a workload: flops_alg = sum(measured operation counts x fixed per-operation
constants, FMA = 2 flop).  The counts come from the -DB200RT_COUNTERS twin of the library (rust_raytracer_b200/
libb200rt_counters.so, built by build.py next to the product library and never on a timed path): this module runs it in a
child process on the requested scene at a few samples per pixel (operation counts per path do not depend on spp).

  python tools/opcount.py book2 800 800 path        # prints the JSON this module returns to bench.py / bench_configs.py
"""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

# SURVEY.md section 8(d).  box: one ray-AABB slab test; sphere / rect / boxprim / tri: one primitive test; a ConstantMedium is
# two boundary tests + 12; finish: HitRecord::new (+ sphere_uv when the texture needs uv); mat: Lambertian, Metal, Dielectric,
# DiffuseLight, Isotropic; tex: Constant, Checker, Image, Perlin turbulence; gen: one primary ray; nee: one light sample.
FLOP = {"box": 25, "sphere": 30, "rect": 12, "boxprim": 30, "tri": 45, "medium_extra": 12, "finish": 35, "sphere_uv": 60, "gen": 40, "nee": 60,
        "mat": [40, 45, 70, 10, 30], "tex": [0, 20, 15, 700]}
INTEGRATORS = {"path": 0, "mixture": 1, "nee": 2}


def _child(name, w, h, integrator, spp):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ctypes as C
    import rust_raytracer_b200 as rt
    from rust_raytracer_b200 import _ffi as F
    from scenes_util import scene_builder
    sb = scene_builder(name)
    ctx = rt.Context(0)
    scene = rt.Scene(ctx, sb)
    p = rt.copy_params(sb.params, width=w, height=h, spp_count=spp, integrator=INTEGRATORS[integrator])
    lib = C.CDLL(F.LIB_PATH)
    out = (C.c_ulonglong * 64)()
    lib.b200rt_debug_counters(out, 1)
    _, st = scene.render(sb.camera, p)
    lib.b200rt_debug_counters(out, 1)
    print(json.dumps({"counts": list(out), "paths": int(st.paths), "extend_rays": int(st.extend_rays), "shadow_rays": int(st.shadow_rays)}))


def algorithmic_flops(name="book2", w=800, h=800, integrator="path", spp=8):
    """-> (flops per path, detail dict) or (None, reason string)."""
    sys.path.insert(0, ROOT)
    from rust_raytracer_b200.build import COUNTERS_LIB
    if not os.path.exists(COUNTERS_LIB):
        return None, "libb200rt_counters.so not built"
    env = dict(os.environ, B200RT_LIB=COUNTERS_LIB, B200RT_WF_PIPES="1")
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", name, str(w), str(h), integrator, str(spp)], env=env,
                       stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    try:
        d = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        return None, "counters run failed: " + r.stderr[-300:]
    c, paths = d["counts"], float(d["paths"])
    # primitive type order of include/b200rt.h: sphere, moving sphere, xy, xz, yz rect, box, triangle, medium sphere, medium box
    prim_flop = [FLOP["sphere"], FLOP["sphere"], FLOP["rect"], FLOP["rect"], FLOP["rect"], FLOP["boxprim"], FLOP["tri"],
                 2 * FLOP["sphere"] + FLOP["medium_extra"], 2 * FLOP["boxprim"] + FLOP["medium_extra"]]
    parts = {
        "box_tests": FLOP["box"] * c[6],
        "primitive_tests": sum(prim_flop[t] * c[10 + t] for t in range(9)),
        "hit_records": FLOP["finish"] * c[7] + FLOP["sphere_uv"] * c[8],
        "scatter": sum(FLOP["mat"][m] * c[40 + m] for m in range(5)),
        "textures": sum(FLOP["tex"][t] * c[48 + t] for t in range(4)),
        "camera": FLOP["gen"] * c[9],
        "light_samples": FLOP["nee"] * c[56],
    }
    per_path = {k: v / paths for k, v in parts.items()}
    rays = float(max(c[0], 1))
    counts = {"rays_per_path": (d["extend_rays"] + d["shadow_rays"]) / paths, "inner_nodes_per_ray": c[1] / rays, "box_tests_per_ray": c[6] / rays, "leaves_per_ray": c[2] / rays,
              "primitive_tests_per_ray": sum(c[10:19]) / rays, "f64_decisions_per_ray": sum(c[20:29]) / rays, "hit_records_per_ray": c[7] / rays}
    return sum(per_path.values()), {"flop_per_path": per_path, "counts": counts, "constants": FLOP}


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--child":
        _child(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), sys.argv[5], int(sys.argv[6]))
    else:
        a = sys.argv[1:] + ["book2", "800", "800", "path"][len(sys.argv) - 1:]
        print(json.dumps(algorithmic_flops(a[0], int(a[1]), int(a[2]), a[3])))
