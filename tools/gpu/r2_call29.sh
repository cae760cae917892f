#!/bin/bash
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
run() { name=$1; shift; env "$@" timeout 400 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > $O/r2_sweep_$name.json 2> $O/r2_sweep_$name.err; python -c "
import json; d=json.load(open('$O/r2_sweep_$name.json')); print('$name', round(d['value'],1), round(d['ms_per_step'],1), 'iters', d['wavefront_iterations_per_step'], 'pipes', d['config']['concurrent_pipelines'])"; }
run base B200RT_X=1
run pipes3 B200RT_WF_PIPES=3
run pipes1 B200RT_WF_PIPES=1
run bins4096 B200RT_WF_NBINS=4096
run batch8 B200RT_WF_BATCH=8
run base2 B200RT_X=1
