#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -4
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1.err
python - <<'P'
import json
d=json.load(open('gpurun_out/bench_n1.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches')}, d['e2e']['value'], d['roofline']['frac'], d['clocks'])
tv=d['tree_verify']; print({k:tv[k] for k in tv if k not in ('qkv_proj','o_proj')}); print(tv.get('qkv_proj')); print(tv.get('o_proj'))
P
# ncu launch list of the same command (serialised, cold cache: compare shares)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"decode_attn|combine_kernel|prep_kernel|linear_|attn_tc|multi_attn|commit_kernel|gemv" -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-graph > gpurun_out/ncu_list.log 2>&1; echo "ncu list exit $?"
CMD="python tools/bench_attn.py --only cfg4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 14 -c 1 -o gpurun_out/prof_attn_tc -f $CMD > gpurun_out/ncu_atc.log 2>&1; echo "ncu attn exit $?"
CMD2="python tools/bench_linear.py --m 512 --layer qkv_proj"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:linear_tc2 -s 10 -c 1 -o gpurun_out/prof_linear_tc2 -f $CMD2 > gpurun_out/ncu_tc2.log 2>&1; echo "ncu tc2 exit $?"
