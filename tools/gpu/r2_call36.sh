#!/bin/bash
# two-ended bins by the dense box: A/B on the mesh config (C5), on a small mesh forced on, and that the book scenes do not take it
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 3 0,2 > $O/r2_split_mesh.jsonl 2> $O/r2_split_mesh.err; echo "mesh rc=$?"
timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 path 2 0,2 > $O/r2_split_mesh_path.jsonl 2>> $O/r2_split_mesh.err; echo "mesh path rc=$?"
timeout 300 python tools/dbg/split_ab.py book2 800 800 256 path 2 0,-1 > $O/r2_split_book2.jsonl 2>> $O/r2_split_mesh.err; echo "book2 rc=$?"
timeout 300 python tools/dbg/split_ab.py book1 1920 1080 64 path 2 0,2 > $O/r2_split_book1.jsonl 2>> $O/r2_split_mesh.err; echo "book1 rc=$?"
tail -3 $O/r2_split_mesh.err
