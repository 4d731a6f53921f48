#!/bin/bash
# TP-shard shapes of the SM-pair GEMM: tile width / split sweep
mkdir -p gpurun_out
( S="512,4096,6144 512,4096,3072 512,4096,1536 512,2048,4096 512,1024,4096 512,512,4096"
  echo "== auto"; FFB_W512=0 timeout 100 python tools/bench_linear_shapes.py $S
  for bn in 128 192 256; do for sp in 1 2; do echo "== BN=$bn S=$sp"; FFB_W512=0 FFB_SK=0 FFB_TC2_BN=$bn FFB_TC2_SPLIT=$sp timeout 100 python tools/bench_linear_shapes.py $S; done; done
) > gpurun_out/ab_tp_shapes.log 2>&1
python - <<'P'
import ast
cur=None; tab={}
for l in open("gpurun_out/ab_tp_shapes.log"):
    if l.startswith("=="): cur=l.strip("= \n"); continue
    if l.startswith("{"):
        d=ast.literal_eval(l); tab.setdefault((d["M"],d["K"],d["N"]),{})[cur]=(round(d["us"],1), round(d["cublas_us"],1))
for k,v in tab.items(): print(k, {c:x[0] for c,x in v.items()}, "cublas", list(v.values())[0][1])
P
