#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
timeout 600 python tools/bench_hit_batch.py > gpurun_out/r2_hit_batch.jsonl 2> gpurun_out/r2_hit_batch.err; echo "rc=$?"; cat gpurun_out/r2_hit_batch.jsonl; tail -n 5 gpurun_out/r2_hit_batch.err
