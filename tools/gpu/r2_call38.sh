#!/bin/bash
# per-ray traversal counts of two extend launches on the mesh config (counters build) for tools/dbg/analyze_rays.py
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
B200RT_WF_SPLIT=0 timeout 300 python tools/dbg/dump_rays.py mesh 2 $O/rays_mesh_2.bin 64 512 > $O/r2_dump_mesh.log 2>&1; echo "dump 2 rc=$?"
B200RT_WF_SPLIT=0 timeout 300 python tools/dbg/dump_rays.py mesh 5 $O/rays_mesh_5.bin 64 512 >> $O/r2_dump_mesh.log 2>&1; echo "dump 5 rc=$?"
tail -4 $O/r2_dump_mesh.log; ls -la $O/*.bin
