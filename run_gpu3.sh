#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 300 > gpurun_out/pytest_gpu_all.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_gpu_all.log
tail -6 gpurun_out/pytest_gpu_all.log
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?" >> gpurun_out/bench.err; cat gpurun_out/bench.json | head -c 1800; tail -2 gpurun_out/bench.err
timeout 300 python tools/bench_linear.py --json gpurun_out/bench_linear.json > gpurun_out/bench_linear.log 2>&1
grep "'M': 16\|'M': 64," gpurun_out/bench_linear.log | cut -c1-160
timeout 600 python tools/bench_attn.py --json gpurun_out/bench_attn.json > gpurun_out/bench_attn.log 2>&1
echo "bench_attn exit $?" >> gpurun_out/bench_attn.log; tail -8 gpurun_out/bench_attn.log
