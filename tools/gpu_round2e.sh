#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_norm_argmax.py tests/test_gpu_cpp_host.py tests/test_gpu_weights_io.py -x -q --timeout 300 2>&1 | tail -5 | tee gpurun_out/t_misc.log
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -x -q --timeout 120 2>&1 | tail -5 | tee gpurun_out/t_lintc.log
timeout 900 python -m pytest tests/test_gpu_linear.py -x -q --timeout 300 2>&1 | tail -5 | tee gpurun_out/t_lin.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/trace_run.log 2> gpurun_out/trace_atc2.txt; echo "trace exit $?"
bash tools/gpu_round2d.sh
timeout 200 python tools/bench_quant.py 2>&1 | tail -12 | tee gpurun_out/bench_quant.log
