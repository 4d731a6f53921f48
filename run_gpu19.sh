#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -m gpu -x -q --timeout 60 2>&1 | tail -15
timeout 300 python tools/bench_linear.py --json gpurun_out/bench_linear_tc2.json 2>&1 | cut -c1-230
