#!/bin/bash
# round 2, final evidence: every config (with CPU baselines and image gates), bench (both arms), launch list, full capture of extend + shade
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 1500 python tools/bench_configs.py --out $O/r2_configs_final.json > $O/r2_configs_final.log 2> $O/r2_configs_final.err; echo "configs rc=$?"
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference_final.json 2> $O/r2_bench_reference_final.err; echo "ref rc=$?"
timeout 600 python bench.py > $O/r2_bench_n1_final.json 2> $O/r2_bench_n1_final.err; echo "bench rc=$?"
timeout 300 python bench.py --workload sppm --steps 4 --warmup 1 > $O/r2_bench_sppm_final.json 2> $O/r2_bench_sppm_final.err; echo "sppm rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches_final.csv python bench.py --scaling weak --spp 128 --steps 2 --warmup 3 --no-cpu-baseline > $O/r2_ncu_launch_final.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 676 --launch-count 2 -o $O/r2_extend_shade_final -f python bench.py --scaling weak --spp 128 --steps 1 --warmup 3 --no-cpu-baseline > $O/r2_ncu_full_final.log 2>&1; echo "ncu full rc=$?"
