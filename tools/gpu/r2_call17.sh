#!/bin/bash
# 8 GPUs: SPPM over 8 devices with the per-context buffer cache (reference configuration and the larger problem)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 200 python tools/bench_sppm.py --no-cpu --gpus 8 --repeat 3 > $O/r2_sppm_8gpu_c17.json 2> $O/r2_sppm_8gpu_c17.err; echo "sppm 8 rc=$?"; cut -c1-400 $O/r2_sppm_8gpu_c17.json
timeout 200 python tools/bench_sppm.py --no-cpu --gpus 8 --photons 4000000 --size 1600 --iterations 10 --spp 16 --repeat 3 > $O/r2_sppm_large_8gpu_c17.json 2> $O/r2_sppm_large_8gpu_c17.err; echo "sppm large 8 rc=$?"; cut -c1-400 $O/r2_sppm_large_8gpu_c17.json
timeout 200 python tools/bench_sppm.py --no-cpu --gpus 4 --repeat 3 > $O/r2_sppm_4gpu_c17.json 2> $O/r2_sppm_4gpu_c17.err; echo "sppm 4 rc=$?"; cut -c1-400 $O/r2_sppm_4gpu_c17.json
