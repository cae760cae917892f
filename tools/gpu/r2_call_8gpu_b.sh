#!/bin/bash
# 8 GPUs: the strong-scaling bench lines of the final build at N = 8, 4, 2 (one box), and the target job as one render_multi call
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
for n in 8 4 2; do
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 5 --warmup 3 > $O/r2_bench_n${n}_final.json 2> $O/r2_bench_n${n}_final.err; echo "bench n$n rc=$?"; head -c 160 $O/r2_bench_n${n}_final.json; echo
done
timeout 300 python tools/full_job.py --scene book2 > $O/r2_full_job_8gpu_final.json 2> $O/r2_full_job_8gpu_final.err; echo "full job rc=$?"; head -c 400 $O/r2_full_job_8gpu_final.json; echo
timeout 300 python tools/full_job.py --scene mesh --check-spp 256 > $O/r2_full_job_mesh_8gpu_final.json 2> $O/r2_full_job_mesh_8gpu_final.err; echo "mesh job rc=$?"; head -c 400 $O/r2_full_job_mesh_8gpu_final.json; echo
