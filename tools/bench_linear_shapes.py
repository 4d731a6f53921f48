"""Projection GEMM at arbitrary shapes: python tools/bench_linear_shapes.py M,K,N [M,K,N ...] (graph replay, rotating weights)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import ops  # noqa: E402
from tools.bench_linear import time_ms  # noqa: E402


def main():
    h = ops.FFHandler("cuda:0")
    for spec in sys.argv[1:]:
        M, K, N = (int(v) for v in spec.split(","))
        nw = max(2, int(300e6 // (2 * K * N)) + 1)
        ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(min(nw, 12))]
        m = ops.LinearMeta(h, K, N, max_tokens_per_batch=1024)
        x = torch.randn(M, K, device="cuda").half()
        out = torch.empty(M, N, device="cuda", dtype=torch.float16)
        i = [0]

        def fn():
            i[0] += 1
            ops.Linear.forward_kernel_wrapper(m, x, out, ws[i[0] % len(ws)], None, K, N, M)
        ms = time_ms(fn, iters=20)

        def cb():
            i[0] += 1
            torch.matmul(x, ws[i[0] % len(ws)].T, out=out)
        cms = time_ms(cb, iters=20)
        print({"M": M, "K": K, "N": N, "us": ms * 1e3, "tflops": 2.0 * M * K * N / ms / 1e9, "cublas_us": cms * 1e3}, flush=True)


if __name__ == "__main__":
    main()
