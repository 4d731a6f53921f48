"""GPU parity: int4/int8 decompression (bit-exact) and the projection GEMMs vs the CPU oracle."""
import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _torch():
    import torch
    return torch


def _gpu_dequant(blob, bits, in_dim, out_dim, dtype="half"):
    torch = _torch()
    from flexflow_b200 import ops
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    out = ops.decompress_weights(torch.from_numpy(blob).cuda(), in_dim, out_dim, B.DT_INT4 if bits == 4 else B.DT_INT8, tdt)
    torch.cuda.synchronize()
    return out.view(torch.int16).cpu().numpy().view(np.uint16)


@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("shape", [(768, 3072), (4096, 1024), (1024, 11008 // 2)])
def test_dequant_bit_exact(bits, shape):
    in_dim, out_dim = shape
    out_dim = (out_dim // 32) * 32
    rng = np.random.default_rng(20)
    w = (rng.standard_normal((out_dim, in_dim)) * 0.02).astype(np.float32)
    blob = O.quantize(w, bits)
    want = O.dequant(blob, bits, in_dim, out_dim)
    got = _gpu_dequant(blob, bits, in_dim, out_dim)
    assert np.array_equal(got, want), int((got != want).sum())
    # and it is a faithful reconstruction
    err = np.abs(want.view(np.float16).astype(np.float32) - w).max()
    assert err < (0.2 / (2 ** bits - 1))


def test_dequant_exhaustive_quotients():
    """Every (q in 0..255) x (positive normal half scale) pair: the GPU's rcp.approx-based __hdiv recipe
    against the oracle's IEEE division (host branch of __hdiv).  Pins the oracle for ALL inputs."""
    # scales: all positive finite halfs >= 2^-14 ... each group column holds one scale; 32 rows hold q values.
    scales = np.arange(0x0400, 0x7C00, dtype=np.uint16)  # 30720 positive normal halfs
    n = len(scales)
    in_dim = n  # one column per scale
    mism = 0
    for qbase in range(0, 256, 32):
        out_dim = 32
        blob = np.zeros(O.blob_bytes(8, in_dim, out_dim), dtype=np.uint8)
        vals = blob[: in_dim * out_dim].reshape(out_dim, in_dim)
        vals[:] = (qbase + np.arange(32, dtype=np.uint8))[:, None]
        meta = blob[in_dim * out_dim:].view(np.uint16)
        meta[:n] = 0x3555  # offset 0.333: exercises the __hadd rounding too
        meta[n:] = scales
        want = O.dequant(blob, 8, in_dim, out_dim)
        got = _gpu_dequant(blob, 8, in_dim, out_dim)
        mism += int((got != want).sum())
    assert mism == 0, mism


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("quant", [0, 8, 4])
@pytest.mark.parametrize("M", [1, 16, 40, 64, 200])
def test_linear_stream(dtype, quant, M, monkeypatch):
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "stream")
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    K, N = 1024, 1536
    rng = np.random.default_rng(21)
    x = O.round_to_dt((rng.standard_normal((M, K))).astype(np.float32), dtype)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    h = ops.FFHandler("cuda:0")
    if quant == 0:
        wq = O.round_to_dt(w, dtype)
        wdev = torch.from_numpy(wq).to("cuda", tdt)
        qt = B.DT_NONE
    else:
        blob = O.quantize(w, quant, dtype)
        bits = O.dequant(blob, quant, K, N, dtype)
        if dtype == "half":
            wq = bits.view(np.float16).astype(np.float32)
        else:
            wq = (bits.astype(np.uint32) << 16).view(np.float32)
        wdev = torch.from_numpy(blob).cuda()
        qt = B.DT_INT8 if quant == 8 else B.DT_INT4
    m = ops.LinearMeta(h, K, N, dtype=tdt, quantization_type=qt, max_tokens_per_batch=256)
    out = torch.empty(M, N, dtype=tdt, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).to("cuda", tdt), out, wdev, None, K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, wq, None, dtype)
    got = out.float().cpu().numpy()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert np.abs(got - want).max() <= tol * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("M,K,N", [(1, 1024, 512), (7, 2048, 1536), (16, 4096, 2048), (16, 4096, 32 * 149)])
def test_linear_gemv_decode(dtype, M, K, N):
    """Default decode path (M <= 16, DT weights): persistent bulk-copy weight stream."""
    torch = _torch()
    from flexflow_b200 import ops
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    rng = np.random.default_rng(23)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), dtype)
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), dtype)
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), dtype)
    h = ops.FFHandler("cuda:0")
    N32 = N  # out_dim % 32 is required by ffb_linear_create
    if N % 32:
        pytest.skip("out_dim must be a multiple of 32")
    m = ops.LinearMeta(h, K, N32, dtype=tdt, max_tokens_per_batch=16)
    out = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
    for _ in range(2):  # twice: the ring/barrier state must be clean across launches
        ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).to("cuda", tdt), out,
                                            torch.from_numpy(w).to("cuda", tdt), torch.from_numpy(b).to("cuda", tdt), K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, w, b, dtype)
    got = out.float().cpu().numpy()
    assert np.isfinite(got).all()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert np.abs(got - want).max() <= tol * max(1.0, np.abs(want).max())


def test_linear_bias_and_errors():
    torch = _torch()
    from flexflow_b200 import ops
    from flexflow_b200.lib import FFBError
    h = ops.FFHandler("cuda:0")
    K, N, M = 256, 64, 3
    rng = np.random.default_rng(22)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.1).astype(np.float32), "half")
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), "half")
    m = ops.LinearMeta(h, K, N, max_tokens_per_batch=8)
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.forward_kernel_wrapper(m, torch.from_numpy(x).cuda().half(), out, torch.from_numpy(w).cuda().half(),
                                      torch.from_numpy(b).cuda().half(), K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, w, b, "half")
    assert np.abs(out.float().cpu().numpy() - want).max() < 2e-2
    with pytest.raises(FFBError):
        ops.Linear.forward_kernel_wrapper(m, torch.from_numpy(x).cuda().half(), out, torch.from_numpy(w).cuda().half(),
                                          None, K, N, 9)
    with pytest.raises(FFBError):
        ops.LinearMeta(h, 100, 64)
