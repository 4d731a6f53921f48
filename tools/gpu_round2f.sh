#!/bin/bash
# round 2, batch f: compact attention roles + evict-first K/V loads + fast division; tcq test; tc4; bench
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn.py tests/test_gpu_attn_ref.py tests/test_gpu_specinfer.py -x -q --timeout 600 2>&1 | tail -4 | tee gpurun_out/f_attn.log
for v in "FFB_ATC_EVICT=1" "FFB_ATC_EVICT=0"; do echo "== $v"; env $v timeout 300 python tools/bench_attn.py --only cfg4 2>&1 | grep "^{"; done | tee gpurun_out/f_attn_evict.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/f_trace_run.log 2> gpurun_out/f_trace_atc.txt; echo "trace exit $?"
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 120 2>&1 | tail -8 | tee gpurun_out/f_lintc.log
FFB_TC4=1 timeout 200 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 60 -k "not tcq" 2>&1 | tail -6 | tee gpurun_out/f_tc4.log
if grep -q "passed" gpurun_out/f_tc4.log && ! grep -q "failed\|error" gpurun_out/f_tc4.log; then
  for v in "FFB_TC4=0" "FFB_TC4=1"; do
    for m in 256 512 1024; do
      echo "== $v M=$m"; env $v timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d['layer'], d['M'], 'ours %.1f us  cublas %.1f  ref16f %.1f' % (d['ms']*1e3, d['cublas_ms']*1e3, d.get('reference_cublas16f_ms',0)*1e3))"
    done
  done | tee gpurun_out/f_tc4_bench.log
  for bn in 128 192 256; do echo "== TC4 BN=$bn"; FFB_TC4=1 FFB_TC4_BN=$bn timeout 120 python tools/bench_linear.py --m 512 2>&1 | grep "^{" | cut -c1-160; done | tee gpurun_out/f_tc4_bn.log
fi
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/f_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['single_launch_bracket']['kernel_ms'], d['layer_split_us'])"
