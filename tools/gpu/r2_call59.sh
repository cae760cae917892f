#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
timeout 50 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "by_default_on_a_large" > gpurun_out/r2_pytest_default_split.log 2>&1; echo "rc=$?"; tail -n 12 gpurun_out/r2_pytest_default_split.log | cut -c1-200
