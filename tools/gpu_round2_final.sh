#!/bin/bash
# end-of-round check of the final build: whole GPU suite, smoke, the bench line, projection micro-benchmark, launch list
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 300 2>&1 | tail -6 > gpurun_out/f_pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/f_smoke.log 2>&1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/f_bench_n1.json 2> gpurun_out/f_bench_n1.err
for m in 64 128 256 512 1024; do timeout 120 python tools/bench_linear.py --m $m 2>&1 | grep "^{"; done > gpurun_out/f_bench_linear.json
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/f_launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-quantized --no-decode > gpurun_out/f_ncu.log 2>&1
cat gpurun_out/f_pytest.log; tail -2 gpurun_out/f_smoke.log
python - <<'P'
import json
d = json.loads(open("gpurun_out/f_bench_n1.json").read().strip().splitlines()[-1])
print(d["value"], d["ms_per_step"], d["e2e"]["value"], d["roofline"]["frac"], d["clocks"], d["layer_split_us"])
P
