#!/bin/bash
# per-lane ray replacement in extend for dense-box scenes (k_wf_extend_dyn, B200RT_WF_DYN=1): mesh config, thresholds and register budgets
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/r2_dyn.jsonl
for t in 16 8 24; do
  echo "{\"lib\": \"shipped (7 blocks)\", \"thresh\": $t}" >> $O/r2_dyn.jsonl
  B200RT_WF_DYN_THRESH=$t timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 0,1 B200RT_WF_DYN >> $O/r2_dyn.jsonl 2>> $O/r2_dyn.err; echo "t=$t rc=$?"
done
for v in dyn8 dyn6; do
  echo "{\"lib\": \"$v\", \"thresh\": 16}" >> $O/r2_dyn.jsonl
  B200RT_LIB=$PWD/tools/dbg/libb200rt_$v.so timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 0,1 B200RT_WF_DYN >> $O/r2_dyn.jsonl 2>> $O/r2_dyn.err; echo "$v rc=$?"
done
cut -c1-250 $O/r2_dyn.jsonl; tail -3 $O/r2_dyn.err
