#!/bin/bash
# 8 GPUs: bench at N=8 and N=4 (strong: the 10,000-spp job), the target job and the C5 mesh job in one b200rt_render_multi call each, SPPM over 8 devices
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r2_bench_n8.json 2> $O/r2_bench_n8.err; echo "bench n8 rc=$?"
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus 4 --steps 5 --warmup 3 > $O/r2_bench_n4.json 2> $O/r2_bench_n4.err; echo "bench n4 rc=$?"
timeout 300 python tools/full_job.py --scene book2 --out $O/r2_book2_10000spp_8gpu.png > $O/r2_full_job_8gpu.json 2> $O/r2_full_job_8gpu.err; echo "full job rc=$?"
timeout 400 python tools/full_job.py --scene mesh --check-spp 256 --out $O/r2_mesh_4096spp_8gpu.png > $O/r2_full_job_mesh_8gpu.json 2> $O/r2_full_job_mesh_8gpu.err; echo "mesh job rc=$?"
timeout 240 python tools/bench_sppm.py --no-cpu --gpus 8 > $O/r2_sppm_8gpu.json 2> $O/r2_sppm_8gpu.err; echo "sppm 8gpu rc=$?"
