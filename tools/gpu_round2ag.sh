#!/bin/bash
# last check of the round: whole GPU suite, smoke, the bench line
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -4 > gpurun_out/h_pytest.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/h_smoke.log 2>&1
timeout 300 python bench.py --steps 20 --warmup 5 --no-quantized > gpurun_out/h_bench_n1.json 2> gpurun_out/h_bench_n1.err
cat gpurun_out/h_pytest.log; tail -1 gpurun_out/h_smoke.log
python - <<'P'
import json
d = json.loads(open("gpurun_out/h_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["layer_split_us"])
P
