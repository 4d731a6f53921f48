"""Phase trace of the SM-pair GEMM (needs linear_tc2.o built with -DFFB_TC2_TRACE_BUILD): python tools/trace_tc2.py M K N"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import ops  # noqa: E402
from flexflow_b200 import lib as ffb_lib  # noqa: E402

M, K, N = (int(x) for x in sys.argv[1:4])
os.environ["FFB_TC2_TRACE"] = "1"
ffb_lib.reload_env()
h = ops.FFHandler("cuda:0")
ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(4)]
x = torch.randn(M, K, device="cuda").half()
out = torch.empty(M, N, device="cuda", dtype=torch.float16)
m = ops.LinearMeta(h, K, N, max_tokens_per_batch=1024)
for i in range(32):  # eager back-to-back launches (the host runs ahead of the device after the first few)
    ops.Linear.forward_kernel_wrapper(m, x, out, ws[i % 4], None, K, N, M)
torch.cuda.synchronize()
