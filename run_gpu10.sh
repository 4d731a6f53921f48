#!/bin/bash
mkdir -p gpurun_out
nvidia-smi -L | head -2
timeout 900 python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -6
timeout 400 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1.err
cat gpurun_out/bench_n1.json | cut -c1-1500
timeout 300 python tools/bench_linear.py --json gpurun_out/bench_linear.json 2>&1 | cut -c1-220
timeout 600 python tools/bench_attn.py --json gpurun_out/bench_attn.json 2>&1 | cut -c1-250
