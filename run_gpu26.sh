#!/bin/bash
timeout 300 python -m pytest tests/test_gpu_linear_tc.py -m gpu -x -q --timeout 60 2>&1 | tail -3
timeout 200 python tools/sweep_tc2.py 512 4096 12288 | grep -E "BN (auto|128|168|192|256)"
timeout 200 python tools/sweep_tc2.py 512 4096 4096 | grep -E "BN (auto|128|256)"
timeout 200 python tools/sweep_tc2.py 1024 4096 12288 | grep -E "BN (auto|168|224|256)"
timeout 200 python tools/sweep_tc2.py 256 4096 12288 | grep -E "BN (auto|168|224|256)"
