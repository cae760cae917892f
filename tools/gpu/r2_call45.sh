#!/bin/bash
# 8 GPUs, final tree (after the builder's size axis and the dense-box bins): the bench line at N=8, the target job and the mesh job as one b200rt_render_multi call each
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29518 bench.py --gpus 8 --steps 5 --warmup 3 > $O/r2_bench_n8_final.json 2> $O/r2_bench_n8_final.err; echo "bench n8 rc=$?"; head -c 160 $O/r2_bench_n8_final.json; echo
timeout 200 python tools/full_job.py --scene book2 > $O/r2_full_job_8gpu_final.json 2> $O/r2_full_job_8gpu_final.err; echo "full job rc=$?"; head -c 400 $O/r2_full_job_8gpu_final.json; echo
timeout 300 python tools/full_job.py --scene mesh --check-spp 256 > $O/r2_full_job_mesh_8gpu_final.json 2> $O/r2_full_job_mesh_8gpu_final.err; echo "mesh job rc=$?"; head -c 400 $O/r2_full_job_mesh_8gpu_final.json; echo
