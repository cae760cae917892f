#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c19.log 2>&1; echo "pytest rc=$?"; tail -n 25 $O/r2_pytest_c19.log | cut -c1-250
