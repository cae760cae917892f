#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 400 python -X faulthandler -m pytest tests -x -q -m gpu --timeout 150 > $O/r2_pytest_2gpu_c16.log 2>&1; echo "pytest 2gpu rc=$?"; tail -n 3 $O/r2_pytest_2gpu_c16.log
for g in 2 1; do
  timeout 300 python tools/bench_sppm.py --no-cpu --gpus $g --repeat 3 > $O/r2_sppm_${g}gpu_c16.json 2> $O/r2_sppm_${g}gpu_c16.err; echo "sppm $g rc=$?"; cut -c1-400 $O/r2_sppm_${g}gpu_c16.json
  timeout 300 python tools/bench_sppm.py --no-cpu --gpus $g --photons 4000000 --size 1600 --iterations 10 --spp 16 --repeat 3 > $O/r2_sppm_large_${g}gpu_c16.json 2> $O/r2_sppm_large_${g}gpu_c16.err; echo "sppm large $g rc=$?"; cut -c1-400 $O/r2_sppm_large_${g}gpu_c16.json
done
