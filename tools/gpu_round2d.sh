#!/bin/bash
# clustered multicast GEMM (linear_tc4, FFB_TC4=1): parity under a short timeout first (a wrong barrier hypothesis hangs),
# then the micro-benchmark with and without it.
mkdir -p gpurun_out
FFB_TC4=1 timeout 120 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 60 2>&1 | tail -6 | tee gpurun_out/t_tc4.log
if grep -q "passed" gpurun_out/t_tc4.log && ! grep -q "failed\|error" gpurun_out/t_tc4.log; then
  for v in "FFB_TC4=0" "FFB_TC4=1"; do
    for m in 256 512 1024; do
      echo "== $v M=$m"; env $v timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d['layer'], d['M'], 'ours %.1f us  cublas %.1f  ref16f %.1f' % (d['ms']*1e3, d['cublas_ms']*1e3, d.get('reference_cublas16f_ms',0)*1e3))"
    done
  done | tee gpurun_out/tc4_bench.log
  for bn in 128 160 192 224 256; do echo "== TC4 BN=$bn"; FFB_TC4=1 FFB_TC4_BN=$bn timeout 120 python tools/bench_linear.py --m 512 2>&1 | grep "^{" | cut -c1-120; done | tee gpurun_out/tc4_bn.log
fi
