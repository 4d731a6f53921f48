#!/bin/bash
# 2 GPUs: TP parity tests, allreduce micro-benchmark (flag-in-data kernel, CTA-count sweep), phase trace, TP=2 bench
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571"
timeout 400 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 300 2>&1 | tail -3 | tee gpurun_out/j_tp2.log
for g in 74 148 296; do echo "== FFB_PEER_LL_GRID=$g"; FFB_PEER_LL_GRID=$g timeout 200 $TR tools/bench_allreduce.py 2>&1 | grep "^{"; done | tee gpurun_out/j_allreduce.log
for mb in 0.25 4; do FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_patrace.so timeout 200 $TR tools/bench_allreduce.py --mb $mb 2>&1 | grep "^{\|PATRACE"; done | tee gpurun_out/j_patrace.log
bash tools/run_tp2_checks.sh 2>&1 | tail -8
