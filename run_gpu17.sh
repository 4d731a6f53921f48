#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_attn.py -m gpu -x -q --timeout 60 2>&1 | tail -5
FFB_ATC_TRACE=40 timeout 60 python tools/bench_attn.py --only cfg4 2> gpurun_out/trace16.txt | tail -1 | cut -c1-120
