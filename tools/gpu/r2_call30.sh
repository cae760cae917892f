#!/bin/bash
# stress: the GPU suite with tiny queue pools (many refills), four concurrent pipelines on every render, with and without the tail kernel
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
i=0
for combo in "B200RT_WF_POOL_LOG2=14 B200RT_WF_PIPES=4 B200RT_WF_PIPE_MIN_LOG2=0" "B200RT_WF_POOL_LOG2=13 B200RT_WF_PIPES=2 B200RT_WF_PIPE_MIN_LOG2=0 B200RT_WF_TAIL=0" "B200RT_WF_POOL_LOG2=16 B200RT_WF_PIPES=3 B200RT_WF_PIPE_MIN_LOG2=0 B200RT_WF_TAIL=100 B200RT_WF_BATCH=1" "B200RT_SPPM_INFLIGHT=1 B200RT_SPPM_PHOTON_ITERS=2" "B200RT_SPPM_INFLIGHT=5 B200RT_SPPM_PHOTON_ITERS=40"; do
  i=$((i+1))
  env $combo timeout 600 python -m pytest tests -x -q -m gpu --timeout 300 > $O/r2_stress_$i.log 2>&1; echo "stress $i [$combo] rc=$?"; tail -n 2 $O/r2_stress_$i.log | cut -c1-200
done
