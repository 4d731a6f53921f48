"""Row operators of SURVEY 8f N1 / N2 beside the reference's own kernels (oracle/_ref/libref_norm.so, compiled from the
reference tree) on the same GPU: us per launch, graph replay over rotating buffers.
    python tools/bench_rowops.py [--json out.json]"""
import ctypes as C
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import ops  # noqa: E402
from oracle import oracle as O  # noqa: E402  (loader of oracle/_ref only)
from tools.bench_linear import time_ms  # noqa: E402


def main():
    ref = O.ref_norm_lib()
    st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)  # noqa: E731
    p = lambda t: C.c_void_p(t.data_ptr())  # noqa: E731
    res, nbuf, hid, vocab = [], 8, 4096, 32000
    for rows in (16, 512):
        a = [torch.randn(rows, hid, device="cuda").half() for _ in range(nbuf)]
        b = [torch.randn(rows, hid, device="cuda").half() for _ in range(nbuf)]
        w = torch.ones(hid, device="cuda").half()
        r, o, nrm = torch.empty_like(a[0]), torch.empty_like(a[0]), torch.empty_like(a[0])
        rms = torch.empty(rows, device="cuda").half()
        m = ops.ResidualRMSNormMeta(None, 1e-5, hid)
        i = [0]

        def ours():
            i[0] += 1
            ops.ResidualRMSNorm.inference_kernel_wrapper(m, None, a[i[0] % nbuf], b[i[0] % nbuf], w, r, o)

        def theirs():
            i[0] += 1
            ref.ref_residual_rms_norm_half(p(a[i[0] % nbuf]), p(b[i[0] % nbuf]), p(w), p(r), p(o), p(rms), p(nrm), rows, hid,
                                           1e-5, st())
        row = dict(op="residual_rms_norm", rows=rows, dim=hid, ours_us=time_ms(ours, iters=40) * 1e3)
        if ref is not None:
            row["reference_kernel_us"] = time_ms(theirs, iters=40) * 1e3
        res.append(row)
        print(row, flush=True)

        x = [torch.randn(rows, vocab, device="cuda").half() for _ in range(nbuf)]
        idx = torch.empty(rows, dtype=torch.int32, device="cuda")
        prob = torch.empty(rows, dtype=torch.float32, device="cuda")
        am = ops.ArgMaxMeta(None)

        def ours_am():
            i[0] += 1
            ops.ArgMax.forward_kernel_wrapper(am, None, x[i[0] % nbuf], idx, None, rows)
        row = dict(op="argmax", rows=rows, length=vocab, ours_us=time_ms(ours_am, iters=40) * 1e3)
        if ref is not None:
            offs = (torch.arange(rows + 1, device="cuda", dtype=torch.int32) * vocab).contiguous()
            kv = torch.zeros(rows * 2, dtype=torch.int64, device="cuda")
            need = C.c_size_t(0)
            ref.ref_argmax_half(p(x[0]), None, None, rows, p(offs), None, 0, C.byref(need), p(kv), st())
            tmp = torch.zeros(max(need.value, 16), dtype=torch.uint8, device="cuda")

            def theirs_am():
                i[0] += 1
                ref.ref_argmax_half(p(x[i[0] % nbuf]), p(idx), p(prob), rows, p(offs), p(tmp), need.value, None, p(kv), st())
            row["reference_cub_us"] = time_ms(theirs_am, iters=40) * 1e3
        res.append(row)
        print(row, flush=True)

        if rows == 16:
            k = 3
            tidx = torch.empty(rows, k, dtype=torch.int32, device="cuda")
            tval = torch.empty(rows, k, dtype=torch.float32, device="cuda")
            tk = ops.ArgTopKMeta(None, k, speculative_decoding=True)

            def ours_tk():
                i[0] += 1
                ops.ArgTopK.forward_kernel_wrapper(tk, None, x[i[0] % nbuf], tval, tidx, rows)

            def theirs_tk():
                i[0] += 1
                ref.ref_arg_topk_half(p(x[i[0] % nbuf]), p(tval), p(tidx), rows, vocab, k, st())
            row = dict(op="arg_topk", rows=rows, length=vocab, k=k, ours_us=time_ms(ours_tk, iters=40) * 1e3)
            if ref is not None:
                row["reference_kernel_us"] = time_ms(theirs_tk, iters=10) * 1e3
            res.append(row)
            print(row, flush=True)
    if "--json" in sys.argv:
        json.dump(res, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)


if __name__ == "__main__":
    main()
