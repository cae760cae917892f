#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 python -m pytest tests/test_sppm.py -x -q -m gpu > $O/r2_pytest_sppm_c6.log 2>&1; echo "pytest sppm rc=$?"; tail -n 3 $O/r2_pytest_sppm_c6.log
for n in 3 5; do
  B200RT_SPPM_INFLIGHT=$n B200RT_SPPM_TRACE=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_c6_inflight$n.json 2> $O/r2_sppm_c6_inflight$n.err; echo "sppm inflight $n rc=$?"
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/r2_sppm_launches_c6.csv python tools/bench_sppm.py --iterations 4 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_launch_c6.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_sppm_gather|k_sppm_knn' --launch-count 4 -o $O/r2_sppm_eye_kernels_c6 -f python tools/bench_sppm.py --iterations 2 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_full_c6.log 2>&1; echo "ncu full rc=$?"
