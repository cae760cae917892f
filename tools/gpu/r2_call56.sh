#!/bin/bash
# after the builder's rewrite (records in build order, parallel halves and axes; same trees): GPU suite, smoke, one bench step
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 200 python -m pytest tests -x -q -m gpu > $O/r2_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_final.log
timeout 60 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 1 $O/r2_smoke_final.log
timeout 60 python tools/dbg/split_ab.py book2 800 800 256 path 1 -1 2>&1 | cut -c1-200
timeout 60 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 1 -1 2>&1 | cut -c1-200
