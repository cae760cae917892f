#!/bin/bash
# 2 GPUs: does SPPM over several devices pay at a larger problem? (4,000,000 photons per iteration, 1600x1600, 10 iterations)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for g in 1 2; do
  timeout 300 python tools/bench_sppm.py --no-cpu --gpus $g --photons 4000000 --size 1600 --iterations 10 --spp 16 --repeat 2 > $O/r2_sppm_large_${g}gpu.json 2> $O/r2_sppm_large_${g}gpu.err; echo "sppm large $g rc=$?"; cut -c1-420 $O/r2_sppm_large_${g}gpu.json
done
