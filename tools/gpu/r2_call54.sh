#!/bin/bash
# 2 GPUs: the whole GPU suite of the final tree (the two-device tests included)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python -X faulthandler -m pytest tests -x -q -m gpu --timeout 120 > $O/r2_pytest_2gpu_final.log 2>&1; echo "pytest 2gpu rc=$?"; tail -n 3 $O/r2_pytest_2gpu_final.log
