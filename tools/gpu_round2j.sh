#!/bin/bash
# 2 GPUs: TP parity tests, allreduce micro-benchmark (protocols, CTA-count sweep), TP=2 bench
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571"
timeout 400 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 300 2>&1 | tail -3 | tee gpurun_out/j_tp2.log
for v in "FFB_PEER_PROTO=l128 FFB_PEER_LL_GRID=148" "FFB_PEER_PROTO=l128 FFB_PEER_LL_GRID=296" "FFB_PEER_PROTO=l128 FFB_PEER_LL_GRID=592" "FFB_PEER_PROTO=ll" "FFB_PEER_PROTO=flag"; do echo "== $v"; env $v timeout 200 $TR tools/bench_allreduce.py 2>&1 | grep "^{" | cut -c1-150; done | tee gpurun_out/j_allreduce.log
bash tools/run_tp2_checks.sh 2>&1 | tail -8
