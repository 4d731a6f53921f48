#!/bin/bash
# round 2, batch i: whole GPU suite, attention Q-first A/B, bench line, round-2 ncu captures
mkdir -p gpurun_out
for v in "FFB_ATC_QFIRST=1" "FFB_ATC_QFIRST=0"; do echo "== $v"; env $v timeout 300 python tools/bench_attn.py --only cfg4 2>&1 | grep "^{" | cut -c1-140; done | tee gpurun_out/i_attn_qfirst.log
FFB_LIB_PATH=$PWD/flexflow_b200/libffb200_atctrace.so FFB_ATC_TRACE=20 timeout 200 python tools/bench_attn.py --only cfg4 > gpurun_out/i_trace_run.log 2> gpurun_out/i_trace_atc.txt; echo "trace exit $?"
timeout 1500 python -m pytest tests -m gpu -x -q --timeout 600 2>&1 | tail -6 | tee gpurun_out/i_all.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/i_bench_n1.json 2> gpurun_out/i_bench_n1.err; echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/i_bench_n1.json')); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['layer_split_us'])"
timeout 300 python tools/bench_linear.py --json gpurun_out/i_bench_linear.json > gpurun_out/i_bench_linear.log 2>&1
timeout 300 python tools/bench_attn.py > gpurun_out/i_bench_attn.log 2>&1
# ncu: launch list of the tree-verify section (eager), then the full captures of the two attention kernels
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"attn_tc|linear_tc2|linear_sk|prep_kernel" -c 640 --csv --log-file gpurun_out/i_launches_tree.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph --no-quantized --no-decode --no-70b > gpurun_out/i_ncu_list.log 2>&1; echo "ncu list exit $?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 14 -c 1 -o gpurun_out/i_attn_tc_full -f python tools/bench_attn.py --only cfg4 > gpurun_out/i_ncu_attn.log 2>&1; echo "ncu attn exit $?"
ncu -i gpurun_out/i_attn_tc_full.ncu-rep --page raw --csv > gpurun_out/i_attn_tc_ncu_full_raw.csv 2>/dev/null
ncu -i gpurun_out/i_attn_tc_full.ncu-rep --page details > gpurun_out/i_attn_tc_ncu_details.txt 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:decode_attn -s 8 -c 1 -o gpurun_out/i_decode_attn_full -f python tools/bench_attn.py --only cfg2 > gpurun_out/i_ncu_dec.log 2>&1; echo "ncu dec exit $?"
ncu -i gpurun_out/i_decode_attn_full.ncu-rep --page raw --csv > gpurun_out/i_decode_attn_ncu_full_raw.csv 2>/dev/null
ncu -i gpurun_out/i_decode_attn_full.ncu-rep --page details > gpurun_out/i_decode_attn_ncu_details.txt 2>/dev/null
timeout 600 ncu --set full --clock-control none --import-source on -k regex:linear_sk -s 10 -c 1 -o gpurun_out/i_linear_sk_full -f python tools/bench_linear.py --m 512 --layer qkv_proj > gpurun_out/i_ncu_sk.log 2>&1; echo "ncu sk exit $?"
ncu -i gpurun_out/i_linear_sk_full.ncu-rep --page raw --csv > gpurun_out/i_linear_sk_ncu_full_raw.csv 2>/dev/null
ncu -i gpurun_out/i_linear_sk_full.ncu-rep --page details > gpurun_out/i_linear_sk_ncu_details.txt 2>/dev/null
ncu -i gpurun_out/i_attn_tc_full.ncu-rep --page source --csv > gpurun_out/i_attn_tc_ncu_source.csv 2>/dev/null
rm -f gpurun_out/*.ncu-rep
ls -la gpurun_out | tail -30
