#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_attn.py -m gpu -x -q --timeout 60 2>&1 | tail -12
for rg in 0 9 18; do echo "== rg $rg"; FFB_ATC_RG=$rg timeout 60 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-120; done
FFB_ATC_DBG=1 timeout 60 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-120
timeout 200 python tools/bench_attn.py 2>&1 | cut -c1-220
