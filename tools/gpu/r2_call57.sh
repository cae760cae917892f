#!/bin/bash
# last build of the round: GPU suite and smoke
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 200 python -m pytest tests -x -q -m gpu > $O/r2_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_final.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 1 $O/r2_smoke_final.log | cut -c1-200
