#!/bin/bash
# Round 2, first GPU call: the SpecInfer closed loop, the whole GPU suite, smoke, the re-headlined bench line.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_specinfer.py -x -q --timeout 600 -s 2>&1 | tail -25 | tee gpurun_out/t_spec.log
timeout 1200 python -m pytest tests -m gpu -x -q --timeout 600 --deselect tests/test_gpu_specinfer.py 2>&1 | tail -6 | tee gpurun_out/t_all.log
timeout 120 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2 | tee gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench exit $?"; tail -3 gpurun_out/bench_n1.err; head -c 1500 gpurun_out/bench_n1.json
nproc; free -g | head -2
