#!/bin/bash
# round 2, fifth GPU call: SPPM with bulk-in-queues + tail photon passes, several passes in flight
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2_pytest_gpu_c5.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_gpu_c5.log
for n in 3 1 2 5; do
  B200RT_SPPM_INFLIGHT=$n B200RT_SPPM_TRACE=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_inflight$n.json 2> $O/r2_sppm_inflight$n.err; echo "sppm inflight $n rc=$?"
done
B200RT_SPPM_PHOTONS_MEGA=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_mega3.json 2> $O/r2_sppm_mega3.err; echo "sppm mega rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/r2_sppm_launches.csv python tools/bench_sppm.py --iterations 4 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_sppm_gather|k_sppm_knn|k_sppm_eye_trace|k_sppm_photons' --launch-count 10 -o $O/r2_sppm_eye_kernels -f python tools/bench_sppm.py --iterations 3 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_full.log 2>&1; echo "ncu full rc=$?"
