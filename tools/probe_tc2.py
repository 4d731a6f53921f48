"""Timing probe of the SM-pair GEMM: K scaling (fixed cost per launch), split-K and weight residency (HBM vs L2).
    python tools/probe_tc2.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import ops  # noqa: E402
from flexflow_b200 import lib as ffb_lib  # noqa: E402
from tools.bench_linear import time_ms  # noqa: E402


def main():
    h = ops.FFHandler("cuda:0")
    cases = []
    for K in (512, 1024, 2048, 4096, 8192):
        cases.append((512, K, 4096, 256, 1))
    for S in (2, 3, 4):
        cases.append((512, 4096, 4096, 256, S))
        cases.append((256, 4096, 4096, 256, S))
    cases.append((256, 4096, 4096, 256, 1))
    cases.append((512, 8192, 4096, 256, 2))
    for M, K, N, bn, S in cases:
        ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(6)]
        x = torch.randn(M, K, device="cuda").half()
        out = torch.empty(M, N, device="cuda", dtype=torch.float16)
        m = ops.LinearMeta(h, K, N, max_tokens_per_batch=1024)
        os.environ["FFB_TC2_BN"] = str(bn)
        ffb_lib.reload_env()
        os.environ["FFB_TC2_SPLIT"] = str(S)
        ffb_lib.reload_env()
        r = []
        for rot in (6, 1):
            i = [0]

            def fn():
                i[0] += 1
                ops.Linear.forward_kernel_wrapper(m, x, out, ws[i[0] % rot], None, K, N, M)
            r.append(min(time_ms(fn) for _ in range(3)) * 1e3)
        print("M %4d K %4d N %4d BN %3d S %d: HBM weights %6.2f us (%4.0f TF)   L2-resident %6.2f us (%4.0f TF)" %
              (M, K, N, bn, S, r[0], 2e-6 * M * K * N / r[0], r[1], 2e-6 * M * K * N / r[1]), flush=True)


if __name__ == "__main__":
    main()
