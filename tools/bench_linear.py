"""Projection GEMM micro-benchmark: TFLOP/s (tcgen05 path) and GB/s (weight-streaming path) at the 7B shapes.
    python tools/bench_linear.py [--json out.json]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import batch_config as B  # noqa: E402
from flexflow_b200 import ops  # noqa: E402


def time_ms(fn, iters=20, warm=3):
    """CUDA-graph replay of `iters` back-to-back calls (removes the Python/ctypes launch cost, which is
    larger than these kernels), timed with CUDA events."""
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def _ref_linear():
    try:
        from oracle import oracle as O  # loader of oracle/_ref only
        return O.ref_linear_lib()
    except Exception:  # noqa: BLE001
        return None


REF_LINEAR = _ref_linear()


def main():
    h = ops.FFHandler("cuda:0")
    res = []
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    only_m = int(sys.argv[sys.argv.index("--m") + 1]) if "--m" in sys.argv else None
    only_layer = sys.argv[sys.argv.index("--layer") + 1] if "--layer" in sys.argv else None
    for name, K, N in (("qkv_proj", 4096, 12288), ("o_proj", 4096, 4096)):
        if only_layer and name != only_layer:
            continue
        ws = [(torch.randn(N, K, device="cuda") * 0.02).half() for _ in range(4)]  # rotate weights: > L2
        for M in ((only_m,) if only_m else (16,) if "--decode-only" in sys.argv else (16, 64, 128, 256, 512, 1024)):
            m = ops.LinearMeta(h, K, N, max_tokens_per_batch=1024)
            x = torch.randn(M, K, device="cuda").half()
            out = torch.empty(M, N, device="cuda", dtype=torch.float16)
            i = [0]

            def fn():
                i[0] += 1
                ops.Linear.forward_kernel_wrapper(m, x, out, ws[i[0] % 4], None, K, N, M)
            ms = time_ms(fn)
            j = [0]

            def ref_fn():  # cuBLAS on the same rotating weights
                j[0] += 1
                torch.matmul(x, ws[j[0] % 4].t(), out=out)
            ref = time_ms(ref_fn)
            flops = 2.0 * M * K * N
            res.append(dict(layer=name, M=M, K=K, N=N, ms=ms, tflops=flops / ms / 1e9, weight_gbs=K * N * 2 / ms / 1e6,
                            cublas_ms=ref, cublas_tflops=flops / ref / 1e9))
            if REF_LINEAR is not None:  # the reference's own call: cublasGemmEx with fp16 accumulation
                import ctypes as C
                k2 = [0]

                def ref16_fn():
                    k2[0] += 1
                    REF_LINEAR.ref_linear_forward_half(C.c_void_p(x.data_ptr()), C.c_void_p(ws[k2[0] % 4].data_ptr()),
                                                       C.c_void_p(out.data_ptr()), K, N, M,
                                                       C.c_void_p(torch.cuda.current_stream().cuda_stream))
                r16 = time_ms(ref16_fn)
                res[-1].update(reference_cublas16f_ms=r16, reference_cublas16f_tflops=flops / r16 / 1e9)
            print(res[-1], flush=True)
    if "--json" in sys.argv:
        json.dump(res, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
