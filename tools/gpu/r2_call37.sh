#!/bin/bash
# size axis in the BVH builder (B200RT_BVH_SIZE_SPLIT) and two-ended bins by the dense box (B200RT_WF_SPLIT): A/B on the mesh config and book-2
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
B200RT_BVH_SIZE_SPLIT=0 timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 0 > $O/r2_split_mesh_bvh0.jsonl 2> $O/r2_split.err; echo "mesh bvh0 rc=$?"
B200RT_BVH_SIZE_SPLIT=1 timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 3 0,2 > $O/r2_split_mesh_bvh1.jsonl 2>> $O/r2_split.err; echo "mesh bvh1 rc=$?"
B200RT_BVH_SIZE_SPLIT=1 timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 64 path 2 0,2 > $O/r2_split_mesh_path_bvh1.jsonl 2>> $O/r2_split.err; echo "mesh path rc=$?"
for r in 1 2; do for v in 0 1; do
B200RT_BVH_SIZE_SPLIT=$v timeout 300 python tools/dbg/split_ab.py book2 800 800 256 path 3 0 > $O/r2_split_book2_bvh${v}_$r.jsonl 2>> $O/r2_split.err; echo "book2 bvh$v rc=$?"
done; done
tail -3 $O/r2_split.err
