#!/bin/bash
# 256 x 512 tile GEMM: parity under a timeout, micro-benchmark on/off, bench
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 100 -k "w512" 2>&1 | tail -8 | tee gpurun_out/v_w512.log
if grep -q "passed" gpurun_out/v_w512.log && ! grep -q "failed\|error" gpurun_out/v_w512.log; then
  for v in "FFB_W512=0" "FFB_W512=auto" "FFB_W512=1"; do
    for m in 256 512 1024; do
      echo "== $v M=$m"; env $v timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d['layer'], d['M'], 'ours %.1f us  cublas %.1f  ref16f %.1f' % (d['ms']*1e3, d['cublas_ms']*1e3, d.get('reference_cublas16f_ms',0)*1e3))"
    done
  done | tee gpurun_out/v_w512_bench.log
  timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --no-quantized --no-decode > gpurun_out/v_bench_n1.json 2> gpurun_out/v_bench_n1.err; python -c "
import json; d=json.load(open('gpurun_out/v_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['layer_split_us'])"
fi
