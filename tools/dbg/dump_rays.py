"""Dump a sample of one extend launch (rays, hits, per-ray node / leaf counts) from a -DB200RT_COUNTERS build.

  python -c "from rust_raytracer_b200.build import build_counters; build_counters()"
  python tools/dbg/dump_rays.py [scene] [iteration] [out.bin] [spp] [warp stride]

The file is read by tools/dbg/analyze_rays.py (offline: lane occupancy of the traversal under alternative queue orders).
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("B200RT_LIB", os.path.join(ROOT, "rust_raytracer_b200", "libb200rt_counters.so"))
name = sys.argv[1] if len(sys.argv) > 1 else "book2"
it = sys.argv[2] if len(sys.argv) > 2 else "12"
out = sys.argv[3] if len(sys.argv) > 3 else os.path.join(ROOT, "gpurun_out", "rays_%s_%s.bin" % (name, it))
spp = int(sys.argv[4]) if len(sys.argv) > 4 else 128
os.environ["B200RT_DBG_DUMP"] = out
os.environ["B200RT_DBG_DUMP_ITER"] = it
os.environ["B200RT_DBG_DUMP_STRIDE"] = sys.argv[5] if len(sys.argv) > 5 else "128"
import rust_raytracer_b200 as rt  # noqa: E402
from scenes_util import scene_builder  # noqa: E402

ctx = rt.Context(0)
sb = scene_builder(name)
sc = rt.Scene(ctx, sb)
p = rt.copy_params(sb.params, spp_count=spp)
acc, st = sc.render(sb.camera, p)
print("scene", name, "paths", st.paths, "extend_rays", st.extend_rays, "iterations", st.wavefront_iterations, "->", out, os.path.getsize(out) if os.path.exists(out) else "MISSING")
