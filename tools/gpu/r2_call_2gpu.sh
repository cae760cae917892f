#!/bin/bash
# 2 GPUs: the whole GPU suite (two-device tests included), bench at N=2 (strong), SPPM over 2 devices and over 1 on the same box
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 400 python -X faulthandler -m pytest tests -x -q -m gpu --timeout 120 > $O/r2_pytest_2gpu.log 2>&1; echo "pytest 2gpu rc=$?"; tail -n 3 $O/r2_pytest_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 > $O/r2_bench_n2.json 2> $O/r2_bench_n2.err; echo "bench n2 rc=$?"; head -c 200 $O/r2_bench_n2.json; echo
timeout 600 python tools/bench_sppm.py --no-cpu --gpus 2 > $O/r2_sppm_2gpu.json 2> $O/r2_sppm_2gpu.err; echo "sppm 2gpu rc=$?"
timeout 600 python tools/bench_sppm.py --no-cpu --gpus 1 > $O/r2_sppm_1gpu_samebox.json 2> $O/r2_sppm_1gpu_samebox.err; echo "sppm 1gpu rc=$?"
