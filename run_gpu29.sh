#!/bin/bash
mkdir -p gpurun_out
for n in 8 4 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540+n)) bench.py --gpus $n --steps 20 --warmup 3 > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n$n exit $?"; tail -2 gpurun_out/bench_n$n.err | cut -c1-300
  python - <<P
import json
d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1])
tv=d.get('tree_verify',{})
print(d['n_gpus'], 'ms', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'e2e', round(d['e2e']['value']), 'attn frac', round(d['roofline']['frac'],3), 'tree ms', tv.get('ms_per_step'), tv.get('error'))
P
  FFB_DISABLE_FUSED_ALLREDUCE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29550+n)) bench.py --gpus $n --steps 20 --warmup 3 --no-tree-verify 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('nccl path:', d['n_gpus'], 'ms', round(d['ms_per_step'],3))"
done
