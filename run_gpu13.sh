#!/bin/bash
mkdir -p gpurun_out
for c in 128 256 384 640; do echo "== chunk $c"; FFB_ATTN_TC_CHUNK=$c timeout 200 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-200; done
CMD="python tools/bench_attn.py --only cfg4"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:attn_tc -s 14 -c 1 -o gpurun_out/prof_atc -f $CMD > gpurun_out/ncu_atc.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_atc.log
