#!/bin/bash
timeout 300 python -m pytest "tests/test_gpu_linear_tc.py::test_linear_tc2_pair_tiles" -m gpu -x -q --timeout 60 -k "1232" 2>&1 | grep -E "assert|Error|error|passed|failed" | head -20
FFB_TC2_BN=128 timeout 60 python tools/check_gemm.py 512 512 1232 512 512 1280 512 512 1216
