#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -m gpu -q --timeout 120 > gpurun_out/pytest_tc.log 2>&1
echo "pytest exit $?" >> gpurun_out/pytest_tc.log
tail -15 gpurun_out/pytest_tc.log
timeout 300 python tools/bench_linear.py --json gpurun_out/bench_linear.json > gpurun_out/bench_linear.log 2>&1
echo "bench_linear exit $?" >> gpurun_out/bench_linear.log
tail -14 gpurun_out/bench_linear.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph"
timeout 300 $CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_attn|combine_kernel|prep_kernel|linear_stream|linear_tc|multi_attn" -c 480 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?" >> gpurun_out/ncu_list.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_attn -s 96 -c 2 -o gpurun_out/prof_decode -f $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full exit $?" >> gpurun_out/ncu_full.log
tail -3 gpurun_out/ncu_full.log
