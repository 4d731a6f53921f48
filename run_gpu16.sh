#!/bin/bash
mkdir -p gpurun_out
for rg in 9 18 16 8 32; do echo "== rg $rg"; FFB_ATC_RG=$rg timeout 60 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-120; done
CMD="python tools/bench_attn.py --only cfg4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 14 -c 1 -o gpurun_out/prof_atc3 -f $CMD > gpurun_out/ncu_atc3.log 2>&1
echo "ncu exit $?"
