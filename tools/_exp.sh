timeout 300 python tools/bench_attn.py --only treeref 2>&1 | tail -3
timeout 400 python tools/bench_attn.py --json gpurun_out/bench_attn.json 2>&1 | grep -c case
