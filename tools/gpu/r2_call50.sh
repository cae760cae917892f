#!/bin/bash
# k_wf_extend_dyn: thresholds 32 (refill only when the whole warp is idle: the static schedule in the dynamic kernel's code) and 1, 4
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/r2_dyn3.jsonl
for t in 32 28 4 1; do
  echo "{\"lib\": \"shipped (7 blocks)\", \"thresh\": $t}" >> $O/r2_dyn3.jsonl
  B200RT_WF_DYN_THRESH=$t timeout 100 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 1 0,1 B200RT_WF_DYN >> $O/r2_dyn3.jsonl 2>> $O/r2_dyn3.err; echo "t=$t rc=$?"
done
cut -c1-250 $O/r2_dyn3.jsonl | grep -v image
