#!/bin/bash
# per-ray traversal counts on the sphere scenes (counters build) for tools/dbg/analyze_rays.py
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python tools/dbg/dump_rays.py book1 2 $O/rays_book1_2.bin 256 64 > $O/r2_dump_b.log 2>&1; echo "rc=$?"
timeout 300 python tools/dbg/dump_rays.py json:scene_500 2 $O/rays_scene500_2.bin 1024 64 >> $O/r2_dump_b.log 2>&1; echo "rc=$?"
timeout 300 python tools/dbg/dump_rays.py cornell_book3 2 $O/rays_cornell_2.bin 64 64 >> $O/r2_dump_b.log 2>&1; echo "rc=$?"
timeout 300 python tools/dbg/dump_rays.py mesh 1 $O/rays_mesh_1.bin 64 1024 >> $O/r2_dump_b.log 2>&1; echo "rc=$?"
tail -5 $O/r2_dump_b.log; ls -la $O/*.bin
