#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > $O/r2_pytest_c14.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_c14.log
timeout 600 python tools/ab.py main=main nocoop=tools/dbg/libb200rt_nocoop.so main2=main nocoop2=tools/dbg/libb200rt_nocoop.so > $O/r2_ab_coop_perlin.txt 2>&1; echo "ab rc=$?"; cat $O/r2_ab_coop_perlin.txt
