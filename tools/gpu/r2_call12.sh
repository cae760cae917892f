#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python -m pytest tests/test_sppm.py -x -q -m gpu --timeout 120 > $O/r2_pytest_c12.log 2>&1; echo "pytest sppm rc=$?"; tail -n 30 $O/r2_pytest_c12.log | cut -c1-240
