#!/bin/bash
# 2-GPU checks: head-sharded TP parity tests and the TP=2 bench line (gpurun --gpus 2 -- tools/run_tp2_checks.sh).
mkdir -p gpurun_out
nvidia-smi -L | head -3
timeout 400 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 300 2>&1 | tail -3 | tee gpurun_out/t_tp2.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err
echo "bench n2 exit $?"; tail -5 gpurun_out/bench_n2.err | cut -c1-400; python -c "
import json; d=json.loads([l for l in open('gpurun_out/bench_n2.json') if l.startswith('{')][-1]); print(d['n_gpus'], 'ms', d['ms_per_step'], 'tok/s', d['value'], 'e2e ms', d['e2e']['ms_per_step'], 'attn', d['roofline']['kernel_ms'], d['roofline']['frac']); print(d.get('tp_parity')); print(d.get('layer_split_us')); print({k: d.get('decode', {}).get(k) for k in ('ms_per_step', 'error')})"
