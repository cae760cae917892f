#!/bin/bash
# ncu full capture of extend + shade on the mesh config (C5) after the size axis and the dense-box bins
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 24 --launch-count 2 -o $O/r2_mesh_split_extend_shade -f python tools/bench_configs.py --no-cpu --only C5 > $O/r2_ncu_mesh_split.log 2>&1; echo "ncu mesh rc=$?"
ls -la $O/*.ncu-rep
