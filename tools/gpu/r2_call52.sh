#!/bin/bash
# 4 GPUs, final tree: the bench line at N=4 and N=2 (strong scaling on the 10,000-spp job)
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for n in 4 2; do
  timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 > $O/r2_bench_n${n}_final.json 2> $O/r2_bench_n${n}_final.err; echo "bench n$n rc=$?"; head -c 160 $O/r2_bench_n${n}_final.json; echo
done
