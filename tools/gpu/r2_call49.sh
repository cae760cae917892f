#!/bin/bash
# k_wf_extend_dyn after the register diet: 7 blocks (shipped build), 8 blocks; and its parity test
mkdir -p gpurun_out; cd "$GRAFT_REPO_ROOT"
O=gpurun_out
: > $O/r2_dyn2.jsonl
timeout 120 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "ray_replacement or two_ended" > $O/r2_pytest_dyn.log 2>&1; echo "pytest rc=$?"; tail -n 3 $O/r2_pytest_dyn.log
echo "{\"lib\": \"shipped (7 blocks)\", \"thresh\": 16}" >> $O/r2_dyn2.jsonl
timeout 100 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 0,1 B200RT_WF_DYN >> $O/r2_dyn2.jsonl 2>> $O/r2_dyn2.err; echo "rc=$?"
echo "{\"lib\": \"dynm8\", \"thresh\": 16}" >> $O/r2_dyn2.jsonl
B200RT_LIB=$PWD/tools/dbg/libb200rt_dynm8.so timeout 100 python tools/dbg/split_ab.py mesh 1920 1080 256 mixture 2 0,1 B200RT_WF_DYN >> $O/r2_dyn2.jsonl 2>> $O/r2_dyn2.err; echo "rc=$?"
cut -c1-250 $O/r2_dyn2.jsonl; tail -3 $O/r2_dyn2.err
