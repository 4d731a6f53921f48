#!/bin/bash
# 2 GPUs: programmatic dependent launch around the allreduce kernel, bench A/B
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581"
for v in 0 1 3; do
  echo "== FFB_AR_PDL=$v"
  FFB_AR_PDL=$v timeout 600 $TR bench.py --gpus 2 --steps 20 --warmup 3 --no-quantized > gpurun_out/o_bench_$v.json 2> gpurun_out/o_bench.err
  python -c "
import json; d=json.loads([l for l in open('gpurun_out/o_bench_$v.json') if l.startswith('{')][-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['layer_split_us']['tree_verify'], d['decode']['ms_per_step'], d['tp_parity']['tree_verify']['rel_err_vs_unsharded'], d['tp_parity']['tree_verify']['ranks_bit_identical'])"
done | tee gpurun_out/o_arpdl.log
FFB_AR_PDL=3 timeout 400 python -m pytest tests/test_gpu_tp.py -m gpu -q --timeout 300 2>&1 | tail -3 | tee gpurun_out/o_tp2.log
