"""GPU parity of ffb_residual_rms_norm / ffb_argmax (csrc/norm_argmax.cu) against the oracle, through the C ABI."""
import ctypes as C

import numpy as np
import pytest

from oracle import oracle as O

pytestmark = pytest.mark.gpu



def _torch():
    import torch
    return torch


def _dev_bits(a):
    torch = _torch()
    return torch.from_numpy(np.ascontiguousarray(a).view(np.int16)).cuda()


def _gpu_norm(x1, x2, w, eps, dt):
    torch = _torch()
    from flexflow_b200 import batch_config as B
    from flexflow_b200 import lib as L
    code = B.DT_HALF if dt == "half" else B.DT_BF16
    d1, dw = _dev_bits(x1), _dev_bits(w)
    d2 = _dev_bits(x2) if x2 is not None else None
    res = torch.zeros_like(d1) if x2 is not None else None
    out = torch.zeros_like(d1)
    p = lambda t: C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)
    rows, n = x1.shape
    L.check(L.load().ffb_residual_rms_norm(p(d1), p(d2), p(dw), p(res), p(out), rows, n, eps, code,
                                           C.c_void_p(torch.cuda.current_stream().cuda_stream)), "rms_norm")
    torch.cuda.synchronize()
    return (None if res is None else res.cpu().numpy().view(np.uint16)), out.cpu().numpy().view(np.uint16)


@pytest.mark.parametrize("dt", ["half", "bf16"])
@pytest.mark.parametrize("n", [64, 768, 4096, 5120, 8192, 16384])
@pytest.mark.parametrize("residual", [False, True])
def test_rms_norm_bit_exact(dt, n, residual):
    """Same roundings and the same fp32 summation order as the reference kernel, so every row is the oracle's bit for
    bit unless rsqrtf's 2 ulp move the 16-bit rms across a rounding boundary (rare: counted, and bounded)."""
    rng = np.random.default_rng(n * 2 + residual)
    rows = 67
    x1 = O.to_bits(rng.standard_normal((rows, n)).astype(np.float32) * 1.5, dt)
    x2 = O.to_bits(rng.standard_normal((rows, n)).astype(np.float32), dt) if residual else None
    w = O.to_bits(1.0 + 0.1 * rng.standard_normal(n).astype(np.float32), dt)
    want_res, want = O.residual_rms_norm(x1, x2, w, 1e-5, dt)
    res, got = _gpu_norm(x1, x2, w, 1e-5, dt)
    if residual:
        assert np.array_equal(res, want_res)
    same = (got == want).all(axis=1)
    assert same.mean() >= 0.97, same.mean()
    np.testing.assert_allclose(O.from_bits(got, dt), O.from_bits(want, dt), rtol=1e-2, atol=1e-6)


def test_rms_norm_large_batch_and_errors():
    torch = _torch()
    from flexflow_b200 import lib as L
    rng = np.random.default_rng(5)
    x1 = O.to_bits(rng.standard_normal((1024, 4096)).astype(np.float32), "half")
    x2 = O.to_bits(rng.standard_normal((1024, 4096)).astype(np.float32), "half")
    w = O.to_bits(np.ones(4096, dtype=np.float32), "half")
    want_res, want = O.residual_rms_norm(x1, x2, w, 1e-6, "half")
    res, got = _gpu_norm(x1, x2, w, 1e-6, "half")
    assert np.array_equal(res, want_res)
    assert (got == want).all(axis=1).mean() >= 0.99
    with pytest.raises(L.FFBError):
        _gpu_norm(x1[:, :40], None, w[:40], 1e-6, "half")  # dim % 32 != 0


@pytest.mark.parametrize("dt", ["half", "bf16"])
@pytest.mark.parametrize("shape", [(16, 32000), (512, 32000), (5, 1001), (3, 50257), (1, 128256)])
def test_argmax_bit_exact(dt, shape):
    torch = _torch()
    from flexflow_b200 import ops
    rng = np.random.default_rng(shape[0])
    x = O.from_bits(O.to_bits(rng.standard_normal(shape).astype(np.float32), dt), dt)
    x[0, 7] = x[0, shape[1] - 1] = 50.0  # tie across threads: smaller index wins
    if shape[0] > 2:
        x[1, :] = 1.0                     # all equal -> 0
        x[2, -1] = 60.0                   # maximum in the scalar tail / last vector
    want_idx, want_prob, _ = O.argmax(x, beam_search=True)
    tdt = torch.float16 if dt == "half" else torch.bfloat16
    d = torch.from_numpy(x).cuda().to(tdt)
    idx = torch.full((shape[0],), -1, dtype=torch.int32, device="cuda")
    parent = torch.full((shape[0],), -1, dtype=torch.int32, device="cuda")
    probs = ops.ArgMax.forward_kernel_wrapper(ops.ArgMaxMeta(None, beam_search=True), None, d, idx, parent, shape[0])
    torch.cuda.synchronize()
    assert np.array_equal(idx.cpu().numpy(), want_idx)
    assert np.array_equal(probs.cpu().numpy(), want_prob)
    assert not parent.cpu().numpy().any()
    idx2 = torch.full((shape[0],), -1, dtype=torch.int32, device="cuda")
    assert ops.ArgMax.forward_kernel_wrapper(ops.ArgMaxMeta(None), None, d, idx2, None, shape[0]) is None
    assert np.array_equal(idx2.cpu().numpy(), want_idx)


def test_norm_wrappers_mirror_reference_names():
    torch = _torch()
    from flexflow_b200 import ops
    rng = np.random.default_rng(9)
    a = torch.from_numpy(rng.standard_normal((16, 4096)).astype(np.float16)).cuda()
    b = torch.from_numpy(rng.standard_normal((16, 4096)).astype(np.float16)).cuda()
    w = torch.from_numpy((1 + 0.1 * rng.standard_normal(4096)).astype(np.float16)).cuda()
    res, out, out2 = torch.empty_like(a), torch.empty_like(a), torch.empty_like(a)
    m = ops.ResidualRMSNormMeta(None, 1e-5, 4096)
    ops.ResidualRMSNorm.inference_kernel_wrapper(m, None, a, b, w, res, out)
    ops.RMSNorm.inference_kernel_wrapper(m, None, res, w, out2)
    torch.cuda.synchronize()
    assert torch.equal(res, a + b)
    assert torch.equal(out, out2)  # norm(residual sum) == residual norm
    want_res, want = O.residual_rms_norm(a.cpu().numpy().view(np.uint16), b.cpu().numpy().view(np.uint16),
                                         w.cpu().numpy().view(np.uint16), 1e-5, "half")
    assert (out.cpu().numpy().view(np.uint16) == want).all(axis=1).mean() >= 0.9


# ---- against the reference's OWN kernels (oracle/_ref/libref_norm.so, compiled from the reference tree) ----
needs_ref = pytest.mark.skipif(O.ref_norm_lib() is None, reason="oracle/_ref/libref_norm.so not built (make -C oracle ref)")


@needs_ref
@pytest.mark.parametrize("n", [64, 768, 4096, 5120, 8192])
@pytest.mark.parametrize("residual", [False, True])
def test_rms_norm_equals_the_reference_kernel_bit_for_bit(n, residual):
    """Same roundings, same summation order, same rsqrtf on the same GPU: every bit of every row."""
    torch = _torch()
    ref = O.ref_norm_lib()
    rng = np.random.default_rng(n + residual)
    rows = 131
    x1 = O.to_bits(rng.standard_normal((rows, n)).astype(np.float32) * 2.0, "half")
    x2 = O.to_bits(rng.standard_normal((rows, n)).astype(np.float32), "half") if residual else None
    w = O.to_bits(1.0 + 0.2 * rng.standard_normal(n).astype(np.float32), "half")
    res, got = _gpu_norm(x1, x2, w, 1e-5, "half")
    d1, dw = _dev_bits(x1), _dev_bits(w)
    out, rms, norm = torch.zeros_like(d1), torch.zeros(rows, dtype=torch.int16, device="cuda"), torch.zeros_like(d1)
    p = lambda t: C.c_void_p(t.data_ptr())
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    if residual:
        d2, rout = _dev_bits(x2), torch.zeros_like(d1)
        assert ref.ref_residual_rms_norm_half(p(d1), p(d2), p(dw), p(rout), p(out), p(rms), p(norm), rows, n, 1e-5, st) == 0
        torch.cuda.synchronize()
        assert np.array_equal(res, rout.cpu().numpy().view(np.uint16))
    else:
        assert ref.ref_rms_norm_half(p(d1), p(dw), p(out), p(rms), p(norm), rows, n, 1e-5, st) == 0
        torch.cuda.synchronize()
    assert np.array_equal(got, out.cpu().numpy().view(np.uint16))


@needs_ref
@pytest.mark.parametrize("shape", [(16, 32000), (64, 32000), (7, 50257)])
def test_argmax_equals_the_reference_cub_call(shape):
    torch = _torch()
    from flexflow_b200 import ops
    ref = O.ref_norm_lib()
    rows, v = shape
    g = torch.Generator(device="cuda").manual_seed(rows)
    x = torch.randn(rows, v, generator=g, device="cuda").half()
    x[0, 11] = x[0, v - 3] = 30.0   # ties resolve to the smaller index in both
    offs = (torch.arange(rows + 1, device="cuda", dtype=torch.int32) * v).contiguous()
    kv = torch.zeros(rows * 2, dtype=torch.int64, device="cuda")
    need = C.c_size_t(0)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert ref.ref_argmax_half(C.c_void_p(x.data_ptr()), None, None, rows, C.c_void_p(offs.data_ptr()), None, 0,
                               C.byref(need), C.c_void_p(kv.data_ptr()), st) == 0
    tmp = torch.zeros(max(need.value, 16), dtype=torch.uint8, device="cuda")
    want_idx = torch.full((rows,), -1, dtype=torch.int32, device="cuda")
    want_prob = torch.zeros(rows, dtype=torch.float32, device="cuda")
    assert ref.ref_argmax_half(C.c_void_p(x.data_ptr()), C.c_void_p(want_idx.data_ptr()), C.c_void_p(want_prob.data_ptr()),
                               rows, C.c_void_p(offs.data_ptr()), C.c_void_p(tmp.data_ptr()), need.value, None,
                               C.c_void_p(kv.data_ptr()), st) == 0
    idx = torch.full((rows,), -1, dtype=torch.int32, device="cuda")
    parent = torch.full((rows,), -1, dtype=torch.int32, device="cuda")
    probs = ops.ArgMax.forward_kernel_wrapper(ops.ArgMaxMeta(None, beam_search=True), None, x, idx, parent, rows)
    torch.cuda.synchronize()
    assert torch.equal(idx, want_idx) and int(idx[0]) == 11
    assert torch.equal(probs, want_prob)


@pytest.mark.parametrize("dt", ["half", "bf16"])
@pytest.mark.parametrize("shape,k", [((16, 32000), 3), ((5, 1001), 2), ((64, 32000), 1), ((3, 50257), 8)])
def test_arg_topk_bit_exact(dt, shape, k):
    torch = _torch()
    from flexflow_b200 import ops
    rng = np.random.default_rng(shape[0] + k)
    x = O.from_bits(O.to_bits(rng.standard_normal(shape).astype(np.float32), dt), dt)
    x[0, [5, shape[1] - 2, 300]] = 40.0   # equal maxima across threads: increasing index order
    x[1, :] = 1.0                         # all equal: 0, 1, 2, ...
    want_idx, want_val = O.arg_topk(x, k)
    tdt = torch.float16 if dt == "half" else torch.bfloat16
    d = torch.from_numpy(x).cuda().to(tdt)
    idx = torch.full((shape[0], k), -1, dtype=torch.int32, device="cuda")
    probs = torch.full((shape[0], k), float("nan"), dtype=torch.float32, device="cuda")
    ops.ArgTopK.forward_kernel_wrapper(ops.ArgTopKMeta(None, k, speculative_decoding=True), None, d, probs, idx, shape[0])
    torch.cuda.synchronize()
    assert np.array_equal(idx.cpu().numpy(), want_idx)
    got_val = probs.cpu().numpy()
    assert np.array_equal(got_val[:, :k - 1], want_val[:, :k - 1])           # the ranks the reference writes
    assert np.array_equal(got_val[:, k - 1], np.take_along_axis(x, want_idx, 1)[:, k - 1])  # and the one it leaves out


@needs_ref
@pytest.mark.parametrize("shape,k", [((16, 32000), 3), ((9, 32000), 2), ((4, 50257), 3)])
def test_arg_topk_equals_the_reference_kernel(shape, k):
    """The reference's own heap kernel (arg_topk.cu, k threads per row) on the same rows, ties included."""
    torch = _torch()
    from flexflow_b200 import ops
    ref = O.ref_norm_lib()
    rows, v = shape
    g = torch.Generator(device="cuda").manual_seed(rows + k)
    x = torch.randn(rows, v, generator=g, device="cuda").half()
    x[0, 11] = x[0, v - 3] = x[0, 500] = 30.0
    x[1, :] = 0.25
    want_idx = torch.full((rows, k), -1, dtype=torch.int32, device="cuda")
    want_val = torch.full((rows, k), float("nan"), dtype=torch.float32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    assert ref.ref_arg_topk_half(C.c_void_p(x.data_ptr()), C.c_void_p(want_val.data_ptr()), C.c_void_p(want_idx.data_ptr()),
                                 rows, v, k, st) == 0
    idx = torch.full((rows, k), -1, dtype=torch.int32, device="cuda")
    probs = torch.full((rows, k), float("nan"), dtype=torch.float32, device="cuda")
    ops.ArgTopK.forward_kernel_wrapper(ops.ArgTopKMeta(None, k, speculative_decoding=True), None, x, probs, idx, rows)
    torch.cuda.synchronize()
    assert torch.equal(idx, want_idx), (idx[:2], want_idx[:2])
    assert torch.equal(probs[:, :k - 1], want_val[:, :k - 1])


@pytest.mark.parametrize("dt", ["half", "bf16"])
@pytest.mark.parametrize("rows,classes", [(16, 32000), (512, 32000), (3, 128256), (5, 1003), (2, 8), (7, 32768)])
def test_softmax_matches_fp32_reference(dt, rows, classes):
    """ffb_softmax (Softmax::inference_kernel_wrapper, the op between lm_head and the token selection): against an fp32
    softmax of the same 16-bit inputs - every probability within one rounding of DT (+ a 2e-6 floor for the ex2 / divide
    approximations), rows sum to 1, the argmax survives, in-place operation allowed."""
    import torch
    from flexflow_b200 import ops
    tdt = torch.float16 if dt == "half" else torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(rows * 7 + classes)
    x = (torch.randn(rows, classes, generator=g, device="cuda") * 3.0).to(tdt)
    if rows > 2:
        x[1] = -1e4 if dt == "half" else -1e30     # a flat row
        x[2, classes // 2] = 60.0                  # one dominant class
    y = torch.empty_like(x)
    ops.Softmax.inference_kernel_wrapper(ops.SoftmaxMeta(None), None, False, x, y)
    torch.cuda.synchronize()
    want = torch.softmax(x.float(), -1)
    eps = 2.0 ** -11 if dt == "half" else 2.0 ** -8
    err = (y.float() - want).abs()
    tol = want * (2 * eps) + (6e-8 if dt == "half" else 1e-30) + 2e-6 * want
    assert bool((err <= tol + 1e-7).all()), float((err - tol).max())
    assert bool(((y.float().sum(-1) - 1).abs() < 2e-2).all())
    top = want.max(-1)
    clear = top.values > want.topk(2, -1).values[:, 1] * (1 + 8 * eps)
    assert bool((y.float().argmax(-1)[clear] == top.indices[clear]).all())
    z = x.clone()
    ops.Softmax.inference_kernel_wrapper(ops.SoftmaxMeta(None), None, False, z, z)
    torch.cuda.synchronize()
    assert torch.equal(z, y)
