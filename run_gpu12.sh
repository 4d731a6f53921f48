#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn.py -m gpu -x -q --timeout 120 2>&1 | tail -25
timeout 300 python tools/bench_attn.py --json gpurun_out/bench_attn_tc.json 2>&1 | cut -c1-250
