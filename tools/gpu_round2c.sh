#!/bin/bash
# attention kernel v2 (4 threads per row): parity tests, micro-benchmark, trace
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attn.py tests/test_gpu_attn_ref.py tests/test_gpu_norm_argmax.py -x -q --timeout 600 2>&1 | tail -8 | tee gpurun_out/t_attn.log
timeout 600 python -m pytest tests/test_gpu_specinfer.py tests/test_gpu_greedy_decode.py -x -q --timeout 600 2>&1 | tail -4 | tee gpurun_out/t_spec.log
for v in "" "FFB_ATC_RG=16"; do
  echo "== $v"; env $v timeout 300 python tools/bench_attn.py 2>&1 | grep "^{" 
done | tee gpurun_out/attn_exp2.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/trace_run.log 2> gpurun_out/trace_atc2.txt; echo "trace exit $?"
