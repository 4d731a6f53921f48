#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_greedy_decode.py -m gpu -x -q --timeout 800 2>&1 | tail -15
