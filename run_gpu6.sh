#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear.py -m gpu -q --timeout 120 -k "gemv or bias" 2>&1 | tail -3
timeout 200 python tools/bench_linear.py --decode-only 2>&1 | cut -c1-190
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['attn_share_of_step'])"
