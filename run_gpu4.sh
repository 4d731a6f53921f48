#!/bin/bash
mkdir -p gpurun_out
for v in "" "FFB_SEPARATE_COMBINE=1" "FFB_DECODE_CHUNK=256" "FFB_DECODE_CHUNK=1024" "FFB_DECODE_U=8" "FFB_DECODE_U=8 FFB_DECODE_CHUNK=1024" "FFB_DECODE_U=2" "FFB_DECODE_U=8 FFB_DECODE_CHUNK=256"; do
  echo "== $v"; env $v timeout 200 python tools/bench_attn.py --only cfg2 2>&1 | tail -1 | cut -c40-200
done
timeout 600 python -m pytest tests/test_gpu_attn.py -m gpu -q --timeout 300 2>&1 | tail -3
