#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -8
timeout 200 python tools/bench_attn.py --json gpurun_out/bench_attn_tc.json 2>&1 | cut -c1-220
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1.err; cut -c1-600 gpurun_out/bench_n1.json
