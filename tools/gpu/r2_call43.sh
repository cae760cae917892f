#!/bin/bash
# round 2, evidence of the final tree (after the builder's size axis): the driver's commands, then the launch list and the full capture of extend + shade on the bench workload
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > $O/r2_pytest_final.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/r2_smoke_final.log 2>&1; echo "smoke rc=$?"; tail -n 2 $O/r2_smoke_final.log
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/r2_bench_reference_final.json 2> $O/r2_bench_reference_final.err; echo "ref rc=$?"
timeout 600 python bench.py > $O/r2_bench_n1_final.json 2> $O/r2_bench_n1_final.err; echo "bench rc=$?"
timeout 300 python bench.py --workload sppm --steps 4 --warmup 1 > $O/r2_bench_sppm_final.json 2> $O/r2_bench_sppm_final.err; echo "sppm rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $O/r2_launches_final.csv python bench.py --scaling weak --spp 128 --steps 2 --warmup 3 --no-cpu-baseline > $O/r2_ncu_launch_final.log 2>&1; echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'k_wf_extend|k_wf_shade' --launch-skip 676 --launch-count 2 -o $O/r2_extend_shade_final -f python bench.py --scaling weak --spp 128 --steps 1 --warmup 3 --no-cpu-baseline > $O/r2_ncu_full_final.log 2>&1; echo "ncu full rc=$?"
