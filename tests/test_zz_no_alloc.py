"""GPU test of ffb_linear_forward_static (the allocation-free variant of the projection GEMM: split-K workspace in a
file-scope __device__ array) against the existing ffb_linear_forward, through the C ABI, on two calls in a row.
Named so that pytest collects it last."""
import ctypes as C

import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as ffb_lib

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


@pytest.mark.parametrize("M,K,N,bias", [(512, 4096, 4096, True), (300, 1024, 1248, False)])
def test_linear_forward_static_matches_linear_forward(M, K, N, bias):
    import torch
    from flexflow_b200 import ops
    L = ffb_lib.load()
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(M, K, generator=g, device="cuda").half()
    w = (torch.randn(N, K, generator=g, device="cuda") * 0.05).half()
    b = torch.randn(N, generator=g, device="cuda").half() if bias else None
    # existing entry point (handle-based)
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=torch.float16, max_tokens_per_batch=1024)
    want = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, x, want, w, b, K, N, M)
    torch.cuda.synchronize()
    assert L.ffb_linear_static_scratch_bytes() <= 32 << 20
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    outs = []
    for _ in range(2):  # two calls in a row share the scratch: the second must not see the first one's leftovers
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        ffb_lib.check(L.ffb_linear_forward_static(_ptr(x), _ptr(out), _ptr(w), _ptr(b) if bias else None, M, K, N,
                                                   B.DT_HALF, stream), "ffb_linear_forward_static")
        torch.cuda.synchronize()
        outs.append(out)
    for out in outs:
        assert torch.isfinite(out.float()).all()
        err = (out.float() - want.float()).abs().max().item()
        assert err <= 2e-3 * max(1.0, want.float().abs().max().item())
    assert torch.equal(outs[0], outs[1])


def test_linear_forward_static_rejects_oversize():
    import torch
    L = ffb_lib.load()
    x = torch.zeros(8, 64, dtype=torch.float16, device="cuda")
    stream = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = L.ffb_linear_forward_static(_ptr(x), _ptr(x), _ptr(x), None, 2048, 4096, 4096, B.DT_HALF, stream)
    assert rc == 1 and b"static-scratch limit" in L.ffb_last_error()
    rc = L.ffb_linear_forward_static(_ptr(x), _ptr(x), _ptr(x), None, 8, 32768, 4096, B.DT_HALF, stream)
    assert rc == 1 and b"static-scratch limits" in L.ffb_last_error()
