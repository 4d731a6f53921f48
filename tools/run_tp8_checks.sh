#!/bin/bash
# 8-GPU checks: TP parity worker at 4 and 8 ranks, allreduce micro-benchmark, bench lines at TP=4 and TP=8
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
for n in 4 8; do
  timeout 300 $TR --nproc-per-node $n --master-port 2955$n tests/tp_worker.py 2>&1 | grep "OK\|rror\|differs\|assert" | head -8 | tee gpurun_out/k_tp$n.log
  timeout 200 $TR --nproc-per-node $n --master-port 2956$n tools/bench_allreduce.py 2>&1 | grep "^{" | tee gpurun_out/k_allreduce$n.log
  timeout 900 $TR --nproc-per-node $n --master-port 2957$n bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n$n exit $?"; tail -3 gpurun_out/bench_n$n.err | cut -c1-300
  python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1]); print(d['n_gpus'], 'ms', d['ms_per_step'], 'tok/s', d['value'], 'e2e ms', d['e2e']['ms_per_step'], 'attn', d['roofline']['kernel_ms'], d['roofline']['frac']); print(d.get('tp_parity')); print(d.get('layer_split_us')); print({k: d.get('decode', {}).get(k) for k in ('ms_per_step', 'error')}); print(d.get('llama70b_tree_verify_tp8'))"
done
