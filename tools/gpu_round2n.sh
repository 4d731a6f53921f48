#!/bin/bash
# programmatic dependent launch of the prep kernel: parity, then bench A/B
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn.py tests/test_gpu_attn_ref.py tests/test_gpu_specinfer.py tests/test_gpu_decode.py -x -q --timeout 300 2>&1 | tail -4 | tee gpurun_out/n_tests.log
for v in "FFB_PREP_PDL=0" "FFB_PREP_PDL=1" "FFB_X=1"; do
  echo "== $v"
  env $v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-quantized > gpurun_out/n_bench_$v.json 2> gpurun_out/n_bench.err
  python -c "
import json; d=json.load(open('gpurun_out/n_bench_$v.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e_graph']['ms_per_step'], d['roofline']['kernel_ms'], d['layer_split_us'], d['decode']['ms_per_step'])"
done | tee gpurun_out/n_pdl.log
