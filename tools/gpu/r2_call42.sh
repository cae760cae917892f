#!/bin/bash
# traversal variants on the mesh config (latency-bound there: 45 MB of scene tables, L1 hit 69 %): stack top in registers, prefetch of the pushed node, 10 blocks per SM, shared short stack
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/r2_mesh_variants.jsonl
for v in shipped topreg nodepf mb10 sstack8 shipped; do
  if [ $v = shipped ]; then unset B200RT_LIB; else export B200RT_LIB=$PWD/tools/dbg/libb200rt_$v.so; fi
  echo "{\"variant\": \"$v\"}" >> $O/r2_mesh_variants.jsonl
  timeout 300 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 2 >> $O/r2_mesh_variants.jsonl 2>> $O/r2_mesh_variants.err; echo "$v rc=$?"
done
unset B200RT_LIB
for v in shipped topreg nodepf; do
  if [ $v = shipped ]; then unset B200RT_LIB; else export B200RT_LIB=$PWD/tools/dbg/libb200rt_$v.so; fi
  echo "{\"variant\": \"$v book2\"}" >> $O/r2_mesh_variants.jsonl
  timeout 300 python tools/dbg/split_ab.py book2 800 800 256 path 2 0 >> $O/r2_mesh_variants.jsonl 2>> $O/r2_mesh_variants.err; echo "$v rc=$?"
done
grep -v image_max $O/r2_mesh_variants.jsonl | cut -c1-200
