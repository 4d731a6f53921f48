timeout 900 python -m pytest tests/test_gpu_linear.py -x -q --timeout 300 2>&1 | tail -5
timeout 200 python tools/bench_quant.py --json gpurun_out/bench_quant.json 2>&1 | grep -v Warn | tail -6
