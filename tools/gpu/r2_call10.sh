#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c10.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/r2_pytest_c10.log
B200RT_SPPM_TRACE=1 timeout 300 python bench.py --workload sppm --steps 4 --warmup 1 > $O/r2_bench_sppm_c10.json 2> $O/r2_bench_sppm_c10.err; echo "sppm bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/r2_sppm_launches_c10.csv python tools/bench_sppm.py --iterations 4 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_launch_c10.log 2>&1; echo "ncu list rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'k_sppm_gather|k_sppm_knn|k_sppm_ph_shade|k_sppm_eye_trace|k_sppm_photons' --launch-skip 4 --launch-count 24 -o $O/r2_sppm_kernels_final -f python tools/bench_sppm.py --iterations 2 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_full_c10.log 2>&1; echo "ncu full rc=$?"
