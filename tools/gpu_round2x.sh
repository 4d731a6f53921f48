#!/bin/bash
mkdir -p gpurun_out
S="512,4096,16896 512,4096,12288 1024,4096,8448"
( echo "== tc2 auto (SK off)"; FFB_SK=0 timeout 120 python tools/bench_linear_shapes.py $S
  echo "== tc4 BN=256"; FFB_TC4=1 FFB_TC4_BN=256 FFB_TC4_DEBUG=1 timeout 120 python tools/bench_linear_shapes.py $S 2>&1 | sort -u
  echo "== tc4 auto"; FFB_TC4=1 FFB_TC4_DEBUG=1 timeout 120 python tools/bench_linear_shapes.py $S 2>&1 | sort -u
  echo "== default"; timeout 120 python tools/bench_linear_shapes.py $S ) > gpurun_out/x_mc.log 2>&1
cat > /tmp/mm.py <<'P'
import torch
x = torch.randn(512, 4096, device="cuda").half(); w = torch.randn(12288, 4096, device="cuda").half(); o = torch.empty(512, 12288, device="cuda", dtype=torch.float16)
for _ in range(3): torch.matmul(x, w.T, out=o)
w2 = torch.randn(4096, 4096, device="cuda").half(); o2 = torch.empty(512, 4096, device="cuda", dtype=torch.float16)
for _ in range(3): torch.matmul(x, w2.T, out=o2)
w3 = torch.randn(1536, 4096, device="cuda").half(); o3 = torch.empty(512, 1536, device="cuda", dtype=torch.float16)
for _ in range(3): torch.matmul(x, w3.T, out=o3)
torch.cuda.synchronize()
P
timeout 200 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__cluster_dim_x,launch__cluster_dim_y,smsp__inst_executed.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none --csv --log-file gpurun_out/x_cublas_names.csv python /tmp/mm.py > gpurun_out/x_cublas_ncu.log 2>&1
tail -30 gpurun_out/x_mc.log; grep -v "^==" gpurun_out/x_cublas_names.csv | cut -c1-400 | tail -70
