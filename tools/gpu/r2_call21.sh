#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c21.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_c21.log | cut -c1-250
B200RT_WF_TAIL=0 timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c21_notail.log 2>&1; echo "pytest (no tail) rc=$?"; tail -n 3 $O/r2_pytest_c21_notail.log | cut -c1-250
timeout 300 python tools/ab.py main=main notail=main,B200RT_WF_TAIL=0 main2=main > $O/r2_ab_tail.txt 2>&1; cat $O/r2_ab_tail.txt
