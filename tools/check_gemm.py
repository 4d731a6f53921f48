"""Spot check of the projection GEMMs at arbitrary shapes against torch.matmul (fp32 accumulate) on the GPU.
    python tools/check_gemm.py M K N [M K N ...]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from flexflow_b200 import ops  # noqa: E402


def main():
    a = [int(x) for x in sys.argv[1:]]
    h = ops.FFHandler("cuda:0")
    g = torch.Generator(device="cuda").manual_seed(0)
    for M, K, N in zip(a[0::3], a[1::3], a[2::3]):
        x = torch.randn(M, K, generator=g, device="cuda", dtype=torch.float16)
        w = (torch.randn(N, K, generator=g, device="cuda", dtype=torch.float32) * K ** -0.5).half()
        m = ops.LinearMeta(h, K, N, dtype=torch.float16, max_tokens_per_batch=1024)
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        worst = 0.0
        for _ in range(5):
            ops.Linear.inference_kernel_wrapper(m, None, x, out, w, None, K, N, M)
            torch.cuda.synchronize()
            want = x.float() @ w.float().t()
            bad = ~torch.isfinite(out)
            err = (out.float() - want).abs().max().item() if not bad.any() else float("nan")
            worst = max(worst, err) if err == err else float("nan")
            if bad.any():
                idx = bad.nonzero()
                print("  non-finite:", int(bad.sum()), "first", idx[0].tolist(), "last", idx[-1].tolist())
                break
        print("M %d K %d N %d  max abs err %s (|want| max %.3f)" % (M, K, N, worst, want.abs().max().item()))
        m.close()


if __name__ == "__main__":
    main()
