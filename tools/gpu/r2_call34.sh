#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
timeout 600 python -X faulthandler -m pytest tests -x -q -m gpu --timeout 200 > gpurun_out/r2_pytest_2gpu_c34.log 2>&1; echo "pytest 2gpu rc=$?"; tail -n 4 gpurun_out/r2_pytest_2gpu_c34.log | cut -c1-200
