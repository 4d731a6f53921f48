#!/bin/bash
mkdir -p gpurun_out
timeout 200 ncu --set full --clock-control none --import-source on -k regex:linear_tcc -s 8 -c 1 -o gpurun_out/ad_tcc -f python tools/bench_linear_shapes.py 64,4096,4096 > gpurun_out/ad_ncu.log 2>&1
ncu -i gpurun_out/ad_tcc.ncu-rep --page details > gpurun_out/ad_tcc_details.txt 2>&1
ncu -i gpurun_out/ad_tcc.ncu-rep --page raw --csv > gpurun_out/ad_tcc_raw.csv 2>&1
grep -i "duration\|throughput\|hit rate\|Cluster\|Grid Size\|SM Frequency" gpurun_out/ad_tcc_details.txt | head -24
