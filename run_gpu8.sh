#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn.py -m gpu -q --timeout 300 2>&1 | tail -4
timeout 600 python tools/bench_attn.py --json gpurun_out/bench_attn.json 2>&1 | cut -c1-250
for c in 256 128 1024; do echo "== chunk $c"; FFB_MULTI_CHUNK=$c timeout 200 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-200; done
