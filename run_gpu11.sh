#!/bin/bash
mkdir -p gpurun_out
CMD="python tools/bench_attn.py --only cfg4"
timeout 200 $CMD > gpurun_out/ba_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:multi_attn -s 14 -c 1 -o gpurun_out/prof_multi -f $CMD > gpurun_out/ncu_multi.log 2>&1
echo "ncu exit $?"; tail -2 gpurun_out/ncu_multi.log
