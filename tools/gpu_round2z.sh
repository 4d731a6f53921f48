#!/bin/bash
mkdir -p gpurun_out
python bench.py --steps 20 --warmup 5 > gpurun_out/z_bench.json 2> gpurun_out/z_bench.err
FFB_W512=0 python bench.py --steps 20 --warmup 5 > gpurun_out/z_bench_w0.json 2> gpurun_out/z_bench_w0.err
python - <<'P'
import json
for f in ("z_bench", "z_bench_w0"):
    try:
        d = json.loads(open("gpurun_out/%s.json" % f).read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d.get("clocks"))
    except Exception as e:
        print(f, "ERR", e)
P
