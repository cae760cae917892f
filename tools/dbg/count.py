"""Per-ray traversal statistics from a -DB200RT_COUNTERS build (tools/dbg/libb200rt_counters.so).

  python -c "from rust_raytracer_b200.build import build_variant; build_variant('counters', ['B200RT_COUNTERS'])"
  B200RT_LIB=tools/dbg/libb200rt_counters.so python tools/dbg/count.py [scene] [pipeline]
"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
os.environ.setdefault("B200RT_LIB", os.path.join(ROOT, "tools", "dbg", "libb200rt_counters.so"))
import rust_raytracer_b200 as rt  # noqa: E402
from rust_raytracer_b200 import _ffi as F  # noqa: E402
from scenes_util import scene_builder  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "book2"
ctx = rt.Context(0)
sb = scene_builder(name)
sc = rt.Scene(ctx, sb)
p = rt.copy_params(sb.params, width=400, height=400, spp_count=8)
lib = C.CDLL(F.LIB_PATH)
out = (C.c_ulonglong * 64)()
lib.b200rt_debug_counters(out, 1)
acc, st = sc.render(sb.camera, p)
lib.b200rt_debug_counters(out, 1)
v = list(out)
rays = max(v[0], 1)
print("scene", name, "rays", v[0], "extend_rays", st.extend_rays, "paths", st.paths)
for i, n in enumerate(["rays", "inner nodes", "leaves", "pushes", "pops culled", "to_object"]):
    print("%-12s %12d  per ray %8.3f" % (n, v[i], v[i] / rays))
types = ["sphere", "moving", "xy", "xz", "yz", "box", "tri", "med_sphere", "med_box"]
print("%-12s %10s %10s %10s   (per ray)" % ("type", "filter", "exact", "accepted"))
for t, n in enumerate(types):
    print("%-12s %10.4f %10.4f %10.4f" % (n, v[10 + t] / rays, v[20 + t] / rays, v[30 + t] / rays))
