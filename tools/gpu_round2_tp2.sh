#!/bin/bash
# 2-GPU check of the final build: TP parity tests + worker, bench line at TP=2
mkdir -p gpurun_out
nvidia-smi -L | wc -l
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
timeout 300 python -m pytest tests/test_gpu_tp.py -x -q --timeout 250 2>&1 | tail -3 | tee gpurun_out/g_tp2_pytest.log
timeout 600 $TR --nproc-per-node 2 --master-port 29572 bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/g_bench_n2.json 2> gpurun_out/g_bench_n2.err
echo "bench n2 exit $?"; tail -3 gpurun_out/g_bench_n2.err | cut -c1-300
python -c "
import json; d=json.loads([l for l in open('gpurun_out/g_bench_n2.json') if l.startswith('{')][-1]); print(d['n_gpus'], 'ms', d['ms_per_step'], 'tok/s', d['value'], 'e2e ms', d['e2e']['ms_per_step'], 'attn', d['roofline']['kernel_ms'], d['roofline']['frac']); print(d.get('tp_parity')); print(d.get('layer_split_us')); print({k: d.get('decode', {}).get(k) for k in ('ms_per_step', 'error')})"
