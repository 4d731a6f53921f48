"""CPU: the oracle's (Residual)RMSNorm and ArgMax restatements (oracle/orc_norm.cpp) against HuggingFace's
LlamaRMSNorm (the alignment path of tests/inference/inference_alignment_test.py) and an independent numpy statement."""
import numpy as np
import pytest

from oracle import oracle as O


def _np_norm(x1, x2, w, eps, dt):
    """Independent statement with numpy float16 arithmetic (half only): every DT store is an astype."""
    xo = x1 if x2 is None else (x1.astype(np.float32) + x2.astype(np.float32)).astype(dt)
    xf = xo.astype(np.float64)
    rms = np.float32(1.0 / np.sqrt((xf * xf).sum(-1, keepdims=True) / x1.shape[-1] + eps)).astype(dt)
    y = (xo.astype(np.float32) * rms.astype(np.float32)).astype(dt)
    return xo, (y.astype(np.float32) * w.astype(np.float32)).astype(dt)


@pytest.mark.parametrize("n", [64, 768, 4096, 5120, 8192])
@pytest.mark.parametrize("residual", [False, True])
def test_rms_norm_against_numpy_half(n, residual):
    rng = np.random.default_rng(n + residual)
    x1 = rng.standard_normal((9, n)).astype(np.float16)
    x2 = rng.standard_normal((9, n)).astype(np.float16) if residual else None
    w = (1.0 + 0.1 * rng.standard_normal(n)).astype(np.float16)
    res, out = O.residual_rms_norm(x1.view(np.uint16), None if x2 is None else x2.view(np.uint16), w.view(np.uint16),
                                   1e-5, "half")
    want_res, want = _np_norm(x1, x2, w, 1e-5, np.float16)
    if residual:
        assert np.array_equal(res, want_res.view(np.uint16))  # a half add has one answer
    # the fp32 summation order differs from numpy's float64 sum: the stored half rms can differ in rare rows
    same_rows = (out == want.view(np.uint16)).all(axis=1)
    assert same_rows.mean() >= 0.8
    np.testing.assert_allclose(O.from_bits(out), want.astype(np.float32), rtol=2e-3, atol=1e-6)


def test_rms_norm_against_huggingface_fp32():
    torch = pytest.importorskip("torch")
    from transformers.models.llama.modeling_llama import LlamaRMSNorm
    torch.manual_seed(0)
    n = 768
    m = LlamaRMSNorm(n, eps=1e-6)
    with torch.no_grad():
        m.weight.copy_(1.0 + 0.1 * torch.randn(n))
        w16 = m.weight.half()
        m.weight.copy_(w16.float())
        a, b = torch.randn(5, n).half(), torch.randn(5, n).half()
        want = m((a + b).float()).numpy()
    res, out = O.residual_rms_norm(a.numpy().view(np.uint16), b.numpy().view(np.uint16), w16.numpy().view(np.uint16),
                                   1e-6, "half")
    assert np.array_equal(res, (a + b).numpy().view(np.uint16))
    np.testing.assert_allclose(O.from_bits(out), want, rtol=3e-3, atol=1e-5)  # three half roundings against fp32


def test_rms_norm_bf16_roundtrip_of_bits():
    rng = np.random.default_rng(3)
    x = rng.standard_normal((4, 256)).astype(np.float32)
    b = O.to_bits(x, "bf16")
    assert np.array_equal(O.to_bits(O.from_bits(b, "bf16"), "bf16"), b)
    w = O.to_bits(np.ones(256, dtype=np.float32), "bf16")
    _, out = O.residual_rms_norm(b, None, w, 1e-5, "bf16")
    xf = O.from_bits(b, "bf16").astype(np.float64)
    want = xf / np.sqrt((xf * xf).mean(-1, keepdims=True) + 1e-5)
    np.testing.assert_allclose(O.from_bits(out, "bf16"), want, rtol=2e-2, atol=1e-6)


def test_argmax_ties_and_beam_outputs():
    rng = np.random.default_rng(0)
    x = rng.standard_normal((7, 32000)).astype(np.float16).astype(np.float32)
    x[2, 100] = x[2, 31999] = 9.0  # tie: the smaller index wins (cub::ArgMax)
    x[3, :] = -np.inf
    idx, prob, parent = O.argmax(x, beam_search=True)
    assert np.array_equal(idx, x.argmax(axis=1).astype(np.int32))
    assert idx[2] == 100 and idx[3] == 0
    assert np.array_equal(prob, x.max(axis=1)) and not parent.any()
    idx2, prob2, parent2 = O.argmax(x)
    assert np.array_equal(idx2, idx) and prob2 is None and parent2 is None


def test_arg_topk_order_and_ties():
    rng = np.random.default_rng(1)
    x = rng.standard_normal((6, 5000)).astype(np.float16).astype(np.float32)
    x[1, [7, 4000, 99]] = 8.0     # three equal maxima: increasing index order
    x[2, :] = 0.5                 # all equal: 0, 1, 2
    idx, val = O.arg_topk(x, 3)
    want = np.argsort(-x, axis=1, kind="stable")[:, :3].astype(np.int32)  # stable: equal values keep index order
    assert np.array_equal(idx, want)
    assert list(idx[1]) == [7, 99, 4000] and list(idx[2]) == [0, 1, 2]
    assert np.array_equal(val[:, :2], np.take_along_axis(x, want, 1)[:, :2]) and np.isnan(val[:, 2]).all()
