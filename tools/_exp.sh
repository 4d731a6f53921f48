timeout 600 python -m pytest tests/test_gpu_attn_ref.py -x -q --timeout 300 2>&1 | tail -8
timeout 300 python tools/bench_attn.py --only cfg2 2>&1 | tail -3
