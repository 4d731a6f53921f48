#!/bin/bash
# One gpurun call: new-kernel tests first, the whole GPU suite, smoke, the bench line, the quantised-projection
# micro-benchmark and a --set full capture of the int8 weight-streaming kernel (each capture only after the same
# command exited 0 without ncu).
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_norm_argmax.py -x -q --timeout 120 2>&1 | tail -5 | tee gpurun_out/t_norm.log
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -4 | tee gpurun_out/t_all.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/smoke.log
timeout 600 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -2 gpurun_out/bench_n1.err
timeout 200 python tools/bench_quant.py --json gpurun_out/bench_quant.json > gpurun_out/bench_quant.log 2>&1; echo "bench_quant exit $?"; cat gpurun_out/bench_quant.log | tail -6
timeout 300 ncu --set full --clock-control none --import-source on -k regex:linear_stream -s 8 -c 1 -o gpurun_out/prof_ls_int8_v2 -f python tools/bench_quant.py --only int8 --layer qkv_proj > gpurun_out/ncu_ls8.log 2>&1; echo "ncu ls8 exit $?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:linear_qgemv -s 8 -c 1 -o gpurun_out/prof_qgemv_oproj_int8 -f python tools/bench_quant.py --only int8 --layer o_proj > gpurun_out/ncu_qg8o.log 2>&1; echo "ncu qgemv exit $?"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"rms_norm|argmax" -c 40 --csv --log-file gpurun_out/launches_norm_argmax.csv python -m pytest tests/test_gpu_norm_argmax.py -x -q > gpurun_out/ncu_norm.log 2>&1; echo "ncu norm exit $?"
ls -la gpurun_out/
