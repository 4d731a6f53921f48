#!/bin/bash
for d in 0 1 2 3; do echo "== dbg $d"; FFB_ATC_DBG=$d timeout 200 python tools/bench_attn.py --only cfg4 2>&1 | tail -1 | cut -c1-120; done
