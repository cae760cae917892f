#!/bin/bash
# full capture of a bulk extend + shade pair of the bench workload (launch 626 of the extend|shade sequence: the second bounce of one pipeline's 41 M paths, 3.6 ms of extend)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 626 --launch-count 2 -o $O/r2_extend_shade_final -f python bench.py --scaling weak --spp 128 --steps 1 --warmup 3 --no-cpu-baseline > $O/r2_ncu_full_final.log 2>&1; echo "ncu full rc=$?"
