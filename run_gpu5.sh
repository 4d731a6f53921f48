#!/bin/bash
mkdir -p gpurun_out
timeout 600 python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/bench.json 2> gpurun_out/bench.err
echo "bench exit $?"; python -c "
import json; d=json.load(open('gpurun_out/bench.json')); print(d['ms_per_step'], d['value'], d['e2e']['ms_per_step'], d['roofline']['kernel_ms'], d['roofline']['frac'], d['roofline']['attn_share_of_step'], d['clocks'])"
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph"
timeout 300 $CMD > gpurun_out/bench_plain.json 2> gpurun_out/bench_plain.err &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_attn|combine_kernel|prep_kernel|linear_stream|linear_tc|multi_attn" -c 384 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list exit $?"
