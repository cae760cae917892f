#!/bin/bash
# ncu full capture of extend + shade on the 318k-triangle mesh config (C5) and of the tail kernel on book-1 (C1)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 24 --launch-count 2 -o $O/r2_mesh_extend_shade -f python tools/bench_configs.py --no-cpu --only C5 > $O/r2_ncu_mesh.log 2>&1; echo "ncu mesh rc=$?"
timeout 600 ncu --set full --clock-control none -k regex:'k_wf_tail' --launch-count 2 -o $O/r2_tail_c1 -f python tools/bench_configs.py --no-cpu --only C1 > $O/r2_ncu_tail.log 2>&1; echo "ncu tail rc=$?"
