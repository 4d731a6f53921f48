#!/bin/bash
# The gpurun command that produced profiles/r01_* (tests, bench, smoke, micro-benchmarks, ncu launch lists and --set full captures).
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -4 | tee gpurun_out/t52.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1.err
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
timeout 300 python tools/bench_attn.py --json gpurun_out/bench_attn.json > gpurun_out/bench_attn.log 2>&1; echo "bench_attn exit $?"
timeout 300 python tools/bench_linear.py --json gpurun_out/bench_linear.json > gpurun_out/bench_linear.log 2>&1; echo "bench_linear exit $?"
# ncu launch lists of the same command (serialised, cold cache: compare shares): decode kernels, then tree-verify kernels
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_attn|combine_kernel|prep_kernel|linear_|gemv" -c 700 --csv --log-file gpurun_out/launches_decode.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-quantized --no-tree-verify > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"attn_tc|linear_tc2|tree_prep|commit|prep_kernel" -s 1600 -c 500 --csv --log-file gpurun_out/launches_tree.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-quantized > gpurun_out/ncu_list2.log 2>&1; echo "ncu list2 exit $?"
CMD="python tools/bench_attn.py --only cfg4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 14 -c 1 -o gpurun_out/prof_attn_tc -f $CMD > gpurun_out/ncu_atc.log 2>&1; echo "ncu attn exit $?"
CMD2="python tools/bench_linear.py --m 512 --layer o_proj"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:linear_tc2 -s 10 -c 1 -o gpurun_out/prof_linear_tc2_oproj -f $CMD2 > gpurun_out/ncu_tc2.log 2>&1; echo "ncu tc2 exit $?"
ls -la gpurun_out/*.ncu-rep gpurun_out/launches_*.csv
