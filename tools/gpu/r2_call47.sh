#!/bin/bash
# book-2 with the box of its 1000-sphere cluster as the dense box (B200RT_BVH_DENSE_PCT=70) and two-ended bins forced on
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
B200RT_BVH_DENSE_PCT=70 timeout 300 python tools/dbg/split_ab.py book2 800 800 256 path 3 0,2 > $O/r2_split_book2_dense70.jsonl 2> $O/r2_split_book2_dense70.err; echo "rc=$?"
cut -c1-230 $O/r2_split_book2_dense70.jsonl
