#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 200 python -m pytest tests/test_sppm.py -x -q -m gpu --timeout 120 > $O/r2_pytest_c18.log 2>&1; echo "pytest sppm rc=$?"; tail -n 2 $O/r2_pytest_c18.log
for v in main gl4 gl16 gl32 main; do
  L=""; if [ $v != main ]; then L=$PWD/tools/dbg/libb200rt_$v.so; fi
  B200RT_LIB=$L timeout 200 python bench.py --workload sppm --steps 4 --warmup 1 --no-cpu-baseline > $O/r2_sppm_c18_$v.json 2> $O/r2_sppm_c18_$v.err; echo "sppm $v rc=$?"
  python -c "
import json; s=json.load(open('$O/r2_sppm_c18_$v.json')); print('$v', round(s['value'],1), {k:round(x,1) for k,x in s['stage_ms'].items() if k!='note'}, round(s['e2e']['value'],1))"
done
