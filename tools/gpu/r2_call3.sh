#!/bin/bash
# round 2, third GPU call: stale-error probe over the GPU suite, SPPM bench of the three-kernel eye pass
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
B200RT_PROBE_ERRORS=1 timeout 1500 python -m pytest tests -q -m gpu > $O/r2_pytest_probe.log 2>&1; echo "probe run rc=$?"; grep -n "probe\]" $O/r2_pytest_probe.log | head; tail -n 4 $O/r2_pytest_probe.log
timeout 600 python -m pytest tests/test_sppm.py -x -q -m gpu > $O/r2_sppm_alone.log 2>&1; echo "sppm alone rc=$?"; tail -n 3 $O/r2_sppm_alone.log
B200RT_SPPM_TRACE=1 timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 > $O/r2_bench_sppm_b.json 2> $O/r2_bench_sppm_b.err; echo "sppm bench rc=$?"
