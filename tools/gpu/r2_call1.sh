#!/bin/bash
# round 2, first GPU call: tests, bench (both arms), launch list, full capture of extend + shade, shade bulk A/B, SPPM bench
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $O/r2_gpu.txt 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu > $O/r2_pytest_gpu.log 2>&1; echo "pytest rc=$?" >> $O/r2_pytest_gpu.log
tail -3 $O/r2_pytest_gpu.log
timeout 600 python bench.py > $O/r2_bench_n1.json 2> $O/r2_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference.json 2> $O/r2_bench_reference.err; echo "ref rc=$?"
timeout 600 python tools/ab.py main=main bulk2=tools/dbg/libb200rt_bulk2.so bulk3=tools/dbg/libb200rt_bulk3.so main2=main > $O/r2_ab_shade_bulk.txt 2>&1; echo "ab rc=$?"
timeout 600 python bench.py --workload sppm --steps 3 --warmup 1 > $O/r2_bench_sppm.json 2> $O/r2_bench_sppm.err; echo "sppm rc=$?"
# launch list (weak step of 128 spp keeps the list short)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches.csv python bench.py --scaling weak --spp 128 --steps 2 --warmup 3 --no-cpu-baseline > $O/r2_ncu_launch.log 2>&1; echo "ncu list rc=$?"
# full capture: the extend + shade launches of one steady-state iteration
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 676 --launch-count 2 -o $O/r2_extend_shade -f python bench.py --scaling weak --spp 128 --steps 1 --warmup 3 --no-cpu-baseline > $O/r2_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la $O | tail -20
