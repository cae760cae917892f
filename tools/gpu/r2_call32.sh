#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c32.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_c32.log | cut -c1-250
timeout 600 python tools/ab.py split=main nosplit=main,B200RT_WF_SPLIT=0 split2=main nosplit2=main,B200RT_WF_SPLIT=0 > $O/r2_ab_split.txt 2>&1; cat $O/r2_ab_split.txt
