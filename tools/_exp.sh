timeout 600 python -m pytest tests/test_gpu_attn_ref.py -x -q --timeout 300 2>&1 | tail -14
