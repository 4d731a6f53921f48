#!/bin/bash
mkdir -p gpurun_out
timeout 500 python -m pytest tests/test_gpu_tp.py -m gpu -x -q --timeout 300 2>&1 | tail -15
for n in 2; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29540+n)) bench.py --gpus $n --steps 20 --warmup 3 --no-tree-verify > gpurun_out/bench_n$n.json 2> gpurun_out/bench_n$n.err
  echo "bench n$n exit $?"; tail -3 gpurun_out/bench_n$n.err | cut -c1-300
  python - <<P
import json
d=json.loads([l for l in open('gpurun_out/bench_n$n.json') if l.startswith('{')][-1])
print(d['n_gpus'], 'ms', round(d['ms_per_step'],3), 'tok/s', round(d['value']), 'e2e', round(d['e2e']['value']))
P
  FFB_DISABLE_FUSED_ALLREDUCE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29550+n)) bench.py --gpus $n --steps 20 --warmup 3 --no-tree-verify 2>/dev/null | python -c "
import json,sys
d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('nccl path:', d['n_gpus'], 'ms', round(d['ms_per_step'],3))"
done
