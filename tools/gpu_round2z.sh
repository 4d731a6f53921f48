#!/bin/bash
mkdir -p gpurun_out
( for e in "A=1" "FFB_PREP_THREADS=256" "FFB_PREP_THREADS=512" "FFB_PREP_BLOCKS_PER_SM=16" "FFB_PREP_THREADS=256 FFB_PREP_BLOCKS_PER_SM=16" "FFB_PREP_THREADS=64 FFB_PREP_BLOCKS_PER_SM=32"; do
    env $e timeout 100 python tools/bench_prep.py
  done
  timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:prep_kernel -c 8 --csv --log-file gpurun_out/z_prep_base.csv python tools/bench_prep.py > /dev/null 2>&1
  FFB_PREP_THREADS=512 timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:prep_kernel -c 8 --csv --log-file gpurun_out/z_prep_512.csv python tools/bench_prep.py > /dev/null 2>&1
  for f in base 512; do echo "ncu $f"; grep prep_kernel gpurun_out/z_prep_$f.csv | awk -F'","' '{print $(NF)}' | tr -d '"' | tr '\n' ' '; echo; done
) > gpurun_out/z_prep.log 2>&1
cat gpurun_out/z_prep.log | cut -c1-250
