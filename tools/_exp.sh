timeout 600 python -m pytest tests/test_gpu_linear.py -x -q -k qgemv --timeout 300 2>&1 | tail -5
timeout 120 python tools/bench_quant.py 2>&1 | grep -o "'layer.*us': [0-9.]*"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:linear_qgemv -s 8 -c 1 -o gpurun_out/prof_qgemv_int8 -f python tools/bench_quant.py --only int8 --layer qkv_proj > gpurun_out/ncu_qg8.log 2>&1; echo "ncu exit $?"
