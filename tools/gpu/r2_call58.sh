#!/bin/bash
# World::new on the mesh config end to end (host compile + upload) after the builder's rewrite
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
timeout 80 python - > gpurun_out/r2_scene_create_mesh.txt 2>&1 <<'PY'
import sys, time, os
sys.path.insert(0, os.getcwd()); sys.path.insert(0, os.path.join(os.getcwd(), "tests"))
import rust_raytracer_b200 as rt
from scenes_util import scene_builder
ctx = rt.Context(0)
sb = scene_builder("mesh"); sb.desc()
for threads in ("1", "16", "16"):
    os.environ["B200RT_BVH_THREADS"] = threads
    t = time.perf_counter(); sc = rt.Scene(ctx, sb); dt = time.perf_counter() - t
    print("World::new, mesh config, %s builder threads: %.1f ms (%.1f MB uploaded)" % (threads, dt * 1e3, sc.info()["h2d_bytes"] / 1e6)); sc.close()
PY
cat gpurun_out/r2_scene_create_mesh.txt
