#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | wc -l
timeout 300 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 200 2>&1 | tail -3
for n in 2 4 8; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540+n)) bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n$n exit $?"; tail -2 gpurun_out/bench_n$n.err | cut -c1-300
  python - <<P
import json
try:
    d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1])
    tv=d.get('tree_verify',{})
    print(d['n_gpus'], 'ms', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'e2e', round(d['e2e']['value']), 'attn frac', round(d['roofline']['frac'],3), 'tree ms', tv.get('ms_per_step'), 'tree attn frac', tv.get('attn_frac_of_hbm_peak'), tv.get('error'))
except Exception as e: print('parse fail', e)
P
done
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 exit $?"
