#!/bin/bash
mkdir -p gpurun_out
( S="1024,1024,8192 1024,8192,1280 512,4096,11008 512,11008,4096"
  echo "== default"; timeout 100 python tools/bench_linear_shapes.py $S
  echo "== wide off"; FFB_W512=0 timeout 100 python tools/bench_linear_shapes.py $S ) > gpurun_out/ae.log 2>&1
cut -c1-200 gpurun_out/ae.log
