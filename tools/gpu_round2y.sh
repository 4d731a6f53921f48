#!/bin/bash
mkdir -p gpurun_out
( timeout 300 python -m pytest tests/test_gpu_linear_tc.py -q -k "w512" 2>&1 | grep -v "^$" | grep "Error\|FAILED\|passed\|failed" | head -20
  S="512,4096,12288 256,4096,32000 384,4096,12288"
  for r in 1 2; do
  echo "== default (8 epilogue warps)"; timeout 120 python tools/bench_linear_shapes.py $S
  echo "== 4 epilogue warps"; FFB_W512_EPI=1 timeout 120 python tools/bench_linear_shapes.py $S
  echo "== wide off"; FFB_W512=0 timeout 120 python tools/bench_linear_shapes.py $S
  done
) > gpurun_out/y_wide.log 2>&1
cut -c1-300 gpurun_out/y_wide.log | tail -40
