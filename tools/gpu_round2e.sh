#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_norm_argmax.py tests/test_gpu_cpp_host.py tests/test_gpu_weights_io.py -x -q --timeout 300 2>&1 | tail -5 | tee gpurun_out/t_misc.log
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 120 2>&1 | tail -5 | tee gpurun_out/t_lintc.log
timeout 900 python -m pytest tests/test_gpu_linear.py -x -q --timeout 300 2>&1 | tail -5 | tee gpurun_out/t_lin.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/trace_run.log 2> gpurun_out/trace_atc2.txt; echo "trace exit $?"
for v in "FFB_ATC_PDL=1" "FFB_ATC_PDL=0"; do echo "== $v"; env $v timeout 300 python tools/bench_attn.py --only cfg4 2>&1 | grep "^{"; done | tee gpurun_out/attn_pdl.log
timeout 600 python -m pytest tests/test_gpu_attn.py tests/test_gpu_attn_ref.py tests/test_gpu_specinfer.py -x -q --timeout 600 2>&1 | tail -4 | tee gpurun_out/t_attn_pdl.log
bash tools/gpu_round2d.sh
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['layer_split_us'])"
timeout 200 python tools/bench_quant.py 2>&1 | tail -12 | tee gpurun_out/bench_quant.log
