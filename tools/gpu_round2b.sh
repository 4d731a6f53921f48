#!/bin/bash
# Round 2, attention experiments: new GPU tests, cfg4 micro-benchmark under three settings, phase trace, launch list.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_weights_io.py -x -q --timeout 300 2>&1 | tail -5 | tee gpurun_out/t_wio.log
for v in "" "FFB_ATC_RG=16" "FFB_ATC_RG=17" "FFB_PREP_CARVEOUT=100" "FFB_PREP_CARVEOUT=50"; do
  echo "== cfg4 $v"; env $v timeout 200 python tools/bench_attn.py --only cfg4 2>&1 | grep -v "^$" | tail -2
done | tee gpurun_out/attn_exp.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/trace_run.log 2> gpurun_out/trace_atc.txt; echo "trace exit $?"; grep -c TRACE gpurun_out/trace_atc.txt
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"attn_tc|linear_tc|prep_kernel|commit" -s 400 -c 300 --csv --log-file gpurun_out/launches_tree.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-quantized --no-decode > gpurun_out/ncu_list2.log 2>&1; echo "ncu list exit $?"
