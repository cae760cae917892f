import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import rust_raytracer_b200._ffi as F
F.LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "libb200rt_dbg.so")
import rust_raytracer_b200 as rt
sys.path.insert(0, "tests")
from scenes_util import scene_builder
import numpy as np
ctx = rt.Context(0)
sb = scene_builder("book2")
sc = rt.Scene(ctx, sb)
p = rt.copy_params(sb.params, width=400, height=400, spp_count=8)
acc, st = sc.render(sb.camera, p)
out = (C.c_ulonglong * 16)()
F.lib().b200rt_debug_counters = F.lib().b200rt_debug_counters
C.CDLL(F.LIB_PATH).b200rt_debug_counters(out, 1)
v = list(out)
rays = v[6]
names = ["classified", "miss", "hit_behind", "hit_accept", "hit_ambiguous", "unsure", "rays", "node_visits"]
print("rays", rays, "extend_rays", st.extend_rays)
for n, x in zip(names, v[:8]): print("%-14s %10d  per ray %.3f" % (n, x, x / rays))
print("unsure by prim type 0..7:", v[8:16])
