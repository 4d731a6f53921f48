#!/bin/bash
# round 2, batch h: attention (one-loop softmax role, Q loads ahead of the K/V stream, L2 prefetch) and stream-K GEMM with L2 prefetch
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_attn.py tests/test_gpu_attn_ref.py tests/test_gpu_specinfer.py -x -q --timeout 600 2>&1 | tail -4 | tee gpurun_out/h_attn.log
for v in "FFB_ATC_PF=0 FFB_ATC_QFIRST=0" "FFB_ATC_PF=0 FFB_ATC_QFIRST=1" "FFB_ATC_PF=4" "FFB_ATC_PF=6" "FFB_ATC_PF=10" "FFB_ATC_PF=16" "FFB_ATC_PF=6 FFB_ATC_EVICT=0"; do echo "== $v"; env $v timeout 300 python tools/bench_attn.py --only cfg4 2>&1 | grep "^{" | cut -c1-140; done | tee gpurun_out/h_attn_sweep.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/h_trace_run.log 2> gpurun_out/h_trace_atc.txt; echo "trace exit $?"
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 100 -k "stream_k" 2>&1 | tail -4 | tee gpurun_out/h_sk.log
for pf in 0 4 8 16 32; do
  for m in 256 512 1024; do
    echo "== FFB_SK=1 PF=$pf M=$m"; FFB_SK=1 FFB_SK_PF=$pf timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{" | python -c "
import sys,ast
for l in sys.stdin:
    d=ast.literal_eval(l); print(d['layer'], d['M'], 'ours %.1f us  cublas %.1f  ref16f %.1f' % (d['ms']*1e3, d['cublas_ms']*1e3, d.get('reference_cublas16f_ms',0)*1e3))"
  done
done | tee gpurun_out/h_sk_bench.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/h_bench_n1.json 2> gpurun_out/h_bench_n1.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/h_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['layer_split_us'])"
