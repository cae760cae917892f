#!/bin/bash
# 1 GPU: gather clipping check, SPPM bench + launch list, every BASELINE config
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python tools/dbg/sppm_clip_check.py > $O/r2_sppm_clip_check.json 2> $O/r2_sppm_clip_check.err; echo "clip check rc=$?"; cat $O/r2_sppm_clip_check.json
timeout 300 python -m pytest tests/test_sppm.py tests/test_multi.py -x -q -m gpu --timeout 120 > $O/r2_pytest_c8.log 2>&1; echo "pytest rc=$?"; tail -n 2 $O/r2_pytest_c8.log
B200RT_SPPM_TRACE=1 timeout 300 python bench.py --workload sppm --steps 3 --warmup 1 > $O/r2_bench_sppm_c8.json 2> $O/r2_bench_sppm_c8.err; echo "sppm bench rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file $O/r2_sppm_launches_c8.csv python tools/bench_sppm.py --iterations 4 --spp 16 --no-cpu --repeat 1 > $O/r2_sppm_ncu_launch_c8.log 2>&1; echo "ncu list rc=$?"
timeout 2400 python tools/bench_configs.py --out $O/r2_configs.json > $O/r2_configs.log 2> $O/r2_configs.err; echo "configs rc=$?"; tail -n 3 $O/r2_configs.err
