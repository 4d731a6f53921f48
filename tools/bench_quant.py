"""Quantised decode projections (BASELINE configs[2]): us per launch and GB/s of blob bytes at M = 16.
    python tools/bench_quant.py [--only int8|int4] [--layer qkv_proj|o_proj] [--json out.json]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import batch_config as B  # noqa: E402
from flexflow_b200 import lib as L  # noqa: E402
from flexflow_b200 import ops  # noqa: E402
from tools.bench_linear import time_ms  # noqa: E402


def make_blob(K, N, bits, gen):
    nvals = K * N * bits // 8
    meta = torch.empty(K * N // 32, dtype=torch.float16, device="cuda")
    return torch.cat([torch.randint(0, 256, (nvals,), generator=gen, device="cuda", dtype=torch.uint8),
                      meta.fill_(-0.04).clone().view(torch.uint8),
                      meta.fill_((2 ** bits - 1) / 0.08).clone().view(torch.uint8)])


def main():
    h = ops.FFHandler("cuda:0")
    only = sys.argv[sys.argv.index("--only") + 1] if "--only" in sys.argv else None
    only_layer = sys.argv[sys.argv.index("--layer") + 1] if "--layer" in sys.argv else None
    M = int(sys.argv[sys.argv.index("--m") + 1]) if "--m" in sys.argv else 16
    gen = torch.Generator(device="cuda").manual_seed(7)
    res = []
    for lname, K, N in (("qkv_proj", 4096, 12288), ("o_proj", 4096, 4096)):
        if only_layer and lname != only_layer:
            continue
        for name, qt, bits in (("int8", B.DT_INT8, 8), ("int4", B.DT_INT4, 4)):
            if only and name != only:
                continue
            nb = 6 if lname == "qkv_proj" else 16  # rotate blobs: > L2
            if os.environ.get("BENCH_QUANT_NB"):
                nb = int(os.environ["BENCH_QUANT_NB"])
            blobs = [make_blob(K, N, bits, gen) for _ in range(nb)]
            m = ops.LinearMeta(h, K, N, torch.float16, qt, max(64, M))
            x = torch.randn(M, K, device="cuda").half()
            out = torch.empty(M, N, device="cuda", dtype=torch.float16)
            i = [0]

            def fn():
                i[0] += 1
                ops.Linear.forward_kernel_wrapper(m, x, out, blobs[i[0] % nb], None, K, N, M)
            ms = time_ms(fn, iters=24)
            res.append(dict(layer=lname, quant=name, M=M, us=ms * 1e3, blob_mb=blobs[0].numel() / 1e6,
                            blob_gbs=blobs[0].numel() / ms / 1e6))
            print(res[-1], flush=True)
            del blobs
    # standalone decompression (what the reference runs before every GEMM when weights are offloaded,
    # src/ops/kernels/linear_kernels.cu:464-490): ours against the reference's own kernel compiled for sm_100a
    if not only and not only_layer:
        import ctypes as C
        from oracle import oracle as O  # the checker's loader of oracle/_ref (bench tools may use it)
        ref = O.ref_lib()
        K, N = 4096, 12288
        for name, qt, bits in (("int8", B.DT_INT8, 8), ("int4", B.DT_INT4, 4)):
            blobs = [make_blob(K, N, bits, gen) for _ in range(4)]
            out = torch.empty(N, K, dtype=torch.float16, device="cuda")
            i = [0]
            fn_ours = (L.load().ffb_dequant_int8 if bits == 8 else L.load().ffb_dequant_int4)

            def ours():
                i[0] += 1
                fn_ours(C.c_void_p(blobs[i[0] % 4].data_ptr()), C.c_void_p(out.data_ptr()), K, N, B.DT_HALF,
                        C.c_void_p(torch.cuda.current_stream().cuda_stream))
            row = dict(layer="qkv_proj dequant only", quant=name, ours_us=time_ms(ours, iters=12) * 1e3,
                       bytes_mb=(blobs[0].numel() + 2 * K * N) / 1e6)
            row["ours_gbs"] = row["bytes_mb"] / row["ours_us"] * 1e3
            if ref is not None:
                fn_ref = ref.ref_decompress_int8_half if bits == 8 else ref.ref_decompress_int4_half

                def theirs():
                    i[0] += 1
                    fn_ref(C.c_void_p(blobs[i[0] % 4].data_ptr()), C.c_void_p(out.data_ptr()), K, N,
                           C.c_void_p(torch.cuda.current_stream().cuda_stream))
                row["reference_kernel_us"] = time_ms(theirs, iters=12) * 1e3
            res.append(row)
            print(row, flush=True)
            del blobs
    if "--json" in sys.argv:
        json.dump(res, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
