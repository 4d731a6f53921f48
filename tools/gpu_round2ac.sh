#!/bin/bash
mkdir -p gpurun_out
( echo "== tc2 short-K rule"; FFB_W512=0 timeout 100 python tools/bench_linear_shapes.py 512,1024,4096 512,2048,4096
  timeout 200 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 90 -k "split_k or pair_tiles" 2>&1 | tail -2
  for rg in 0 16 17 18 19; do echo "== attn FFB_ATC_RG=$rg"; FFB_ATC_RG=$rg timeout 120 python tools/bench_prep.py; done
) > gpurun_out/ac.log 2>&1
cut -c1-220 gpurun_out/ac.log
