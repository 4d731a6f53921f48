timeout 600 python -m pytest tests/test_gpu_linear.py tests/test_gpu_norm_argmax.py -x -q -k "qgemv or argmax" --timeout 300 2>&1 | tail -3
echo "== qgemv forced"; FFB_LINEAR=qgemv timeout 120 python tools/bench_quant.py 2>&1 | grep -o "'layer.*us': [0-9.]*"
echo "== default"; timeout 120 python tools/bench_quant.py 2>&1 | grep -o "'layer.*us': [0-9.]*"
timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-tree-verify --no-quantized | python -c "import json,sys; print(json.dumps(json.load(sys.stdin)['block_ops'],indent=0))"
