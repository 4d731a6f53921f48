#!/bin/bash
mkdir -p gpurun_out
for bn in 128 256; do
CMD="python tools/bench_linear.py --m 512 --layer qkv_proj"
FFB_TC2_BN=$bn timeout 100 $CMD 2>&1 | tail -1 | cut -c1-200
FFB_TC2_BN=$bn timeout 300 ncu --set full --clock-control none --import-source on -k regex:linear_tc2 -s 10 -c 1 -o gpurun_out/prof_tc2_bn$bn -f $CMD > gpurun_out/ncu_tc2_$bn.log 2>&1
echo "ncu exit $?"
done
