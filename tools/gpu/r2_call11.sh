#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python -m pytest tests/test_sppm.py -x -q -m gpu --timeout 200 > $O/r2_pytest_c11.log 2>&1; echo "pytest sppm rc=$?"; tail -n 30 $O/r2_pytest_c11.log | cut -c1-240
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_sppm_gather|k_sppm_knn' --launch-count 6 -o $O/r2_sppm_eye_final -f python tools/bench_sppm.py --iterations 3 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_full_c11.log 2>&1; echo "ncu full rc=$?"
