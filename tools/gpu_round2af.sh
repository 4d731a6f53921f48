#!/bin/bash
# cluster split-K kernel above 128 tokens on narrow layers: parity under a timeout, then the shapes it takes, on / off
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 90 -k "tcc" 2>&1 | grep -v "^$" | grep "Error\|FAILED\|passed\|failed\|assert\|Timeout" | head -12 | tee gpurun_out/af_test.log
if grep -q "passed" gpurun_out/af_test.log && ! grep -q "failed\|rror" gpurun_out/af_test.log; then
  S="512,4096,1536 256,4096,4096 256,4096,2048 512,4096,1024 384,4096,1536"
  ( echo "== default"; timeout 100 python tools/bench_linear_shapes.py $S
    echo "== FFB_TCC=0"; FFB_TCC=0 timeout 100 python tools/bench_linear_shapes.py $S ) 2>&1 | cut -c1-160 | tee gpurun_out/af_bench.log
fi
