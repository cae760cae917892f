#!/bin/bash
# validation of the final tree: the driver's GPU commands (reference arm and SPPM line as in r2_call43.sh, unchanged since)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python -m pytest tests -x -q -m gpu > $O/r2_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_final.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2_smoke_final.log
timeout 300 python bench.py --steps 3 --warmup 3 > $O/r2_bench_n1_final.json 2> $O/r2_bench_n1_final.err; echo "bench rc=$?"; head -c 300 $O/r2_bench_n1_final.json; echo
