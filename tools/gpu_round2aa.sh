#!/bin/bash
# cluster split-K kernel for <= 128 tokens: parity under a timeout, then the micro-benchmark on / off
mkdir -p gpurun_out
timeout 400 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 90 -k "tcc or test_linear_tc_fp" 2>&1 | grep -v "^$" | grep "Error\|FAILED\|passed\|failed\|assert\|Timeout" | head -20 | tee gpurun_out/aa_test.log
if grep -q "passed" gpurun_out/aa_test.log && ! grep -q "failed\|rror" gpurun_out/aa_test.log; then
  fmt='
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d["layer"], d["M"], "ours %.1f us  cublas %.1f  ref16f %.1f" % (d["ms"]*1e3, d["cublas_ms"]*1e3, d.get("reference_cublas16f_ms",0)*1e3))'
  for v in "A=1" "FFB_TCC=0"; do
    for m in 17 32 64 128; do
      echo "== $v M=$m"; env $v timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "$fmt"
    done
  done | tee gpurun_out/aa_bench.log
fi
