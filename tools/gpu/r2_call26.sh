#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for p in 26 28 29 30; do
  B200RT_WF_POOL_LOG2=$p timeout 400 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r2_strong_pool$p.json 2> $O/r2_strong_pool$p.err; echo "pool $p rc=$?"
  python -c "
import json; d=json.load(open('$O/r2_strong_pool$p.json')); print('pool 2^$p', round(d['value'],1), round(d['ms_per_step'],1), 'e2e', round(d['e2e']['value'],1), d['e2e']['ms_per_step_list'], 'iters', d['wavefront_iterations_per_step'])"
done
