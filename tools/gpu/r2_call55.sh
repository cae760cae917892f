#!/bin/bash
# host side only: stage times of compile_scene on the mesh config with 1, 4, 16 builder threads (the GPU box's host cores)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
nproc > $O/r2_compile_timing.txt
for t in 1 4 16; do echo "threads $t" >> $O/r2_compile_timing.txt; B200RT_BVH_THREADS=$t B200RT_COMPILE_TIMING=1 timeout 100 python tools/dbg/compile_timing.py >> $O/r2_compile_timing.txt 2>&1; done
tail -40 $O/r2_compile_timing.txt
