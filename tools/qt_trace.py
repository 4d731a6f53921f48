import sys, os, torch
sys.path.insert(0, "/root/repo")
from flexflow_b200 import batch_config as B, ops
sys.path.insert(0, "/root/repo/tools")
from bench_quant import make_blob
h = ops.FFHandler("cuda:0")
gen = torch.Generator(device="cuda").manual_seed(7)
for K, N in ((4096, 4096), (4096, 12288)):
    blobs = [make_blob(K, N, 8, gen) for _ in range(4)]
    m = ops.LinearMeta(h, K, N, torch.float16, B.DT_INT8, 64)
    x = torch.randn(16, K, device="cuda").half(); out = torch.empty(16, N, device="cuda", dtype=torch.float16)
    print("== N", N, file=sys.stderr)
    for i in range(8):
        ops.Linear.forward_kernel_wrapper(m, x, out, blobs[i % 4], None, K, N, 16)
    torch.cuda.synchronize()
