#!/bin/bash
# round 2, second GPU call: the stale-error hunt in test_sppm, the remaining GPU tests, compute-sanitizer memcheck + racecheck
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
export B200RT_LIB=$PWD/tools/dbg/libb200rt_head.so
timeout 600 python -m pytest tests/test_sppm.py -x -q -m gpu > $O/r2_dbg_sppm_alone.log 2>&1; echo "sppm alone rc=$?"
timeout 900 python -m pytest tests -q -m gpu > $O/r2_pytest_gpu_all.log 2>&1; echo "all (no -x) rc=$?"; tail -5 $O/r2_pytest_gpu_all.log
timeout 900 compute-sanitizer --tool memcheck --report-api-errors all --show-backtrace host --log-file $O/r2_dbg_api_errors.log python -m pytest tests/test_multi.py tests/test_sppm.py -x -q -m gpu > $O/r2_dbg_api_errors.out 2>&1; echo "api errors rc=$?"
timeout 1200 compute-sanitizer --tool memcheck --log-file $O/r2_memcheck.log python tools/gpu/sanitize_target.py > $O/r2_memcheck.out 2>&1; echo "memcheck rc=$?"; tail -3 $O/r2_memcheck.log
timeout 1500 compute-sanitizer --tool racecheck --log-file $O/r2_racecheck.log python tools/gpu/sanitize_target.py > $O/r2_racecheck.out 2>&1; echo "racecheck rc=$?"; tail -3 $O/r2_racecheck.log
