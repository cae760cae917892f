"""Host only: wall time of compile_scene (World::new without the upload) on the mesh config, three times.

  B200RT_COMPILE_TIMING=1 B200RT_BVH_THREADS=16 python tools/dbg/compile_timing.py   # stage times on stderr
"""
import sys, time
import os; ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))); sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import rust_raytracer_b200 as rt
from scenes_util import scene_builder
sb = scene_builder("mesh"); d = sb.desc()
for _ in range(3):
    t=time.time(); rt.compile_stats(d); print("compile %.3f s"%(time.time()-t))
