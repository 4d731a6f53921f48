#!/bin/bash
# round 2, batch g: stream-K projection GEMM (linear_sk.cu): parity under a timeout, then the micro-benchmark on/off
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 100 -k "stream_k" 2>&1 | tail -8 | tee gpurun_out/g_sk.log
if grep -q "passed" gpurun_out/g_sk.log && ! grep -q "failed\|error" gpurun_out/g_sk.log; then
  timeout 300 python -m pytest tests/test_gpu_linear_tc.py tests/test_gpu_linear.py -x -q --timeout 100 2>&1 | tail -4 | tee gpurun_out/g_lin_all.log
  for v in "FFB_SK=0" "FFB_SK=auto" "FFB_SK=1"; do
    for m in 192 256 512 1024; do
      echo "== $v M=$m"; env $v timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d['layer'], d['M'], 'ours %.1f us  cublas %.1f  ref16f %.1f' % (d['ms']*1e3, d['cublas_ms']*1e3, d.get('reference_cublas16f_ms',0)*1e3))"
    done
  done | tee gpurun_out/g_sk_bench.log
  for p in 64 72 74; do echo "== SK pairs=$p"; FFB_SK=1 FFB_SK_PAIRS=$p timeout 120 python tools/bench_linear.py --m 512 2>&1 | grep "^{" | cut -c1-110; done | tee gpurun_out/g_sk_pairs.log
  timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/g_bench_n1.json 2> gpurun_out/g_bench_n1.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/g_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['layer_split_us'])"
fi
