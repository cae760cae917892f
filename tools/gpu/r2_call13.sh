#!/bin/bash
# round 2, validation of the final tree: the driver's commands
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > $O/r2_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2_smoke_final.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference_final.json 2> $O/r2_bench_reference_final.err; echo "ref rc=$?"
timeout 600 python bench.py > $O/r2_bench_n1_final.json 2> $O/r2_bench_n1_final.err; echo "bench rc=$?"
timeout 300 python bench.py --workload sppm --steps 4 --warmup 1 > $O/r2_bench_sppm_final.json 2> $O/r2_bench_sppm_final.err; echo "sppm rc=$?"
