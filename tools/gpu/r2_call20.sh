#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 200 > $O/r2_pytest_c20.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_c20.log | cut -c1-250
for t in 4096 0; do
  B200RT_WF_TAIL=$t timeout 600 python tools/bench_configs.py --no-cpu --only C1,C2,C2n,C4 --out $O/r2_configs_tail$t.json > $O/r2_configs_tail$t.log 2>&1; echo "configs tail=$t rc=$?"
  python -c "
import json
for r in json.load(open('$O/r2_configs_tail$t.json'))['rows']: print('tail=$t', r['config'], round(r['mpaths_s'],1), round(r['ms_per_call'],2), r['pipelines'])"
done
