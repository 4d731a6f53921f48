"""In-process sweep of the SM-pair GEMM's tile width (FFB_TC2_BN) at one shape: python tools/sweep_tc2.py M K N"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import ops  # noqa: E402
from flexflow_b200 import lib as ffb_lib  # noqa: E402
from tools.bench_linear import time_ms  # noqa: E402


def main():
    M, K, N = (int(x) for x in sys.argv[1:4])
    h = ops.FFHandler("cuda:0")
    ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(4)]
    x = torch.randn(M, K, device="cuda").half()
    out = torch.empty(M, N, device="cuda", dtype=torch.float16)
    m = ops.LinearMeta(h, K, N, max_tokens_per_batch=1024)
    for bn in ["auto", 96, 112, 128, 144, 160, 168, 176, 192, 208, 224, 240, 256]:
        if bn == "auto":
            os.environ.pop("FFB_TC2_BN", None)
            ffb_lib.reload_env()
        else:
            os.environ["FFB_TC2_BN"] = str(bn)
            ffb_lib.reload_env()
        i = [0]

        def fn():
            i[0] += 1
            ops.Linear.forward_kernel_wrapper(m, x, out, ws[i[0] % 4], None, K, N, M)
        ms = min(time_ms(fn) for _ in range(3))
        print("M %d K %d N %d BN %s: %.2f us  %.0f TFLOP/s" % (M, K, N, bn, ms * 1e3, 2.0 * M * K * N / ms / 1e9), flush=True)


if __name__ == "__main__":
    main()
