#!/bin/bash
# 1 GPU: every BASELINE config with FP32 / HBM fractions, CPU baseline and the image gate
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 2400 python tools/bench_configs.py --out $O/r2_configs.json > $O/r2_configs.log 2> $O/r2_configs.err; echo "configs rc=$?"; tail -n 3 $O/r2_configs.err
