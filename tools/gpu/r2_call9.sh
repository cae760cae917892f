#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for n in 3 5 7; do
  B200RT_SPPM_INFLIGHT=$n timeout 200 python bench.py --workload sppm --steps 3 --warmup 1 --no-cpu-baseline > $O/r2_sppm_c9_inflight$n.json 2> $O/r2_sppm_c9_inflight$n.err; echo "sppm inflight $n rc=$?"
done
