#!/bin/bash
# round 2, fourth GPU call: the driver's GPU test command twice, SPPM with the wavefront photon pass (A/B against the round-1 photon kernel), launch list + full capture of the new SPPM kernels
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for r in 1 2; do timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2_pytest_gpu_run$r.log 2>&1; echo "pytest run $r rc=$?"; tail -n 2 $O/r2_pytest_gpu_run$r.log; done
B200RT_SPPM_TRACE=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_wf.json 2> $O/r2_sppm_wf.err; echo "sppm wf rc=$?"
B200RT_SPPM_PHOTONS_MEGA=1 B200RT_SPPM_TRACE=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_mega.json 2> $O/r2_sppm_mega.err; echo "sppm mega rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file $O/r2_sppm_launches.csv python tools/bench_sppm.py --iterations 4 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_launch.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_sppm_gather|k_sppm_knn|k_sppm_ph_shade|k_sppm_eye_trace' --launch-count 12 -o $O/r2_sppm_kernels -f python tools/bench_sppm.py --iterations 2 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_full.log 2>&1; echo "ncu full rc=$?"
