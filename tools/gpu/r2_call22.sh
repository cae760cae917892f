#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for t in 4096 16384 65536 262144; do
  B200RT_WF_TAIL=$t timeout 300 python tools/bench_configs.py --no-cpu --only C1,C2,C3a,C5 --out $O/r2_configs_tail$t.json > $O/r2_configs_tail$t.log 2>&1
  python -c "
import json
print('tail=$t', ' '.join('%s %.0f' % (r['config'], r['mpaths_s']) for r in json.load(open('$O/r2_configs_tail$t.json'))['rows']))"
done
timeout 400 python tools/ab.py t4k=main,B200RT_WF_TAIL=4096 t16k=main,B200RT_WF_TAIL=16384 t64k=main,B200RT_WF_TAIL=65536 t256k=main,B200RT_WF_TAIL=262144 > $O/r2_ab_tail_sweep.txt 2>&1; cat $O/r2_ab_tail_sweep.txt
