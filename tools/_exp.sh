timeout 600 python -m pytest tests/test_gpu_norm_argmax.py -x -q -k "reference" --timeout 300 2>&1 | tail -12
