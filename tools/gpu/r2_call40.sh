#!/bin/bash
# after the builder's size axis and the dense-box bins: GPU tests, all configs re-measured, the bench line
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu > $O/r2_pytest_split.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_split.log
timeout 900 python tools/bench_configs.py --no-cpu --out $O/r2_configs_split.json > $O/r2_configs_split.log 2>&1; echo "configs rc=$?"
timeout 600 python bench.py > $O/r2_bench_n1_split.json 2> $O/r2_bench_n1_split.err; echo "bench rc=$?"; cat $O/r2_bench_n1_split.json | cut -c1-400
