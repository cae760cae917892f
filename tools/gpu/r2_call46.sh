#!/bin/bash
# leaf-against-split with a traversal cost in the builder (B200RT_BVH_TRAV_COST): mesh config and book-2
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/r2_trav_cost.jsonl
for v in 0 0.5 1 2 0; do
  echo "{\"trav_cost\": \"$v\", \"scene\": \"mesh\"}" >> $O/r2_trav_cost.jsonl
  B200RT_BVH_TRAV_COST=$v timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 -1 >> $O/r2_trav_cost.jsonl 2>> $O/r2_trav_cost.err; echo "mesh $v rc=$?"
done
for v in 0 0.5 1 2 0; do
  echo "{\"trav_cost\": \"$v\", \"scene\": \"book2\"}" >> $O/r2_trav_cost.jsonl
  B200RT_BVH_TRAV_COST=$v timeout 300 python tools/dbg/split_ab.py book2 800 800 256 path 2 -1 >> $O/r2_trav_cost.jsonl 2>> $O/r2_trav_cost.err; echo "book2 $v rc=$?"
done
grep -v image_max $O/r2_trav_cost.jsonl | cut -c1-175
