#!/bin/bash
# multicast experiment: SM-pair tiles (linear_tc2) against clusters of two pairs sharing A by TMA multicast (linear_tc4) at whole-wave shapes
mkdir -p gpurun_out
S="512,4096,18944 512,4096,12288 1024,4096,9472"
( echo "== tc2 BN=256"; FFB_SK=0 FFB_TC2_BN=256 FFB_TC2_SPLIT=1 timeout 120 python tools/bench_linear_shapes.py $S
  echo "== tc4 BN=256"; FFB_TC4=1 FFB_TC4_BN=256 FFB_TC4_DEBUG=1 timeout 120 python tools/bench_linear_shapes.py $S 2>&1 | sort -u
  echo "== tc4 auto"; FFB_TC4=1 FFB_TC4_DEBUG=1 timeout 120 python tools/bench_linear_shapes.py $S 2>&1 | sort -u
  echo "== default"; timeout 120 python tools/bench_linear_shapes.py $S ) > gpurun_out/w_mc.log 2>&1
tail -40 gpurun_out/w_mc.log
