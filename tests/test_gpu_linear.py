"""GPU parity: int4/int8 decompression (bit-exact) and the projection GEMMs vs the CPU oracle."""
import ctypes as C

import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as ffb_lib  # noqa: E402
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
    """Every (q in 0..255) x (finite non-zero half scale: both signs, subnormals included) pair: the GPU's
    rcp.approx-based __hdiv recipe (two-at-a-time form, csrc/dequant.cuh) against the oracle's IEEE division (host
    branch of __hdiv).  Pins the oracle - and the dead correction branch - for ALL inputs."""
    pos = np.arange(0x0001, 0x7C00, dtype=np.uint16)
    scales = np.concatenate([pos, pos | 0x8000, pos[:2]])  # 63488 = multiple of 8; one column per scale
    n = len(scales)
    in_dim = n
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


def test_dequant_degenerate_scales_take_the_elementwise_rule():
    """scale = 0 / inf / NaN columns (never produced by the quantiser): the pair rule needs a finite reciprocal, so
    these fall back to the element-wise __hdiv recipe; finite results must still match, the rest must be non-finite."""
    in_dim, out_dim = 64, 32
    blob = np.zeros(O.blob_bytes(8, in_dim, out_dim), dtype=np.uint8)
    blob[: in_dim * out_dim] = np.arange(in_dim * out_dim, dtype=np.uint32).astype(np.uint8)
    meta = blob[in_dim * out_dim:].view(np.uint16)
    meta[:in_dim] = 0x3555
    meta[in_dim:] = 0x4D00  # 20.0
    meta[in_dim + 3], meta[in_dim + 10], meta[in_dim + 17] = 0x0000, 0x7C00, 0x7E00
    want = O.dequant(blob, 8, in_dim, out_dim).view(np.float16).astype(np.float32)
    got = _gpu_dequant(blob, 8, in_dim, out_dim).view(np.float16).astype(np.float32)
    fin = np.isfinite(want)
    assert (got[fin] == want[fin]).all() and not np.isfinite(got[~fin]).any()


@pytest.mark.parametrize("quant", [8, 4])
def test_linear_stream_dequantises_bit_exactly(quant, monkeypatch):
    """The weight-streaming GEMM dequantises in registers (byte / nibble extraction of its own): with one-hot
    activations row m of the output IS column k0+m of the dequantised matrix, so it can be compared bit for bit."""
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "stream")
    ffb_lib.reload_env()
    K, N = 256, 96
    rng = np.random.default_rng(5)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant)
    want = O.dequant(blob, quant, K, N)  # [N][K] half bits
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4, max_tokens_per_batch=16)
    bdev = torch.from_numpy(blob).cuda()
    for k0 in range(0, K, 16):
        x = torch.zeros(16, K, dtype=torch.float16, device="cuda")
        x[torch.arange(16), k0 + torch.arange(16)] = 1.0
        out = torch.empty(16, N, dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, out, bdev, None, K, N, 16)
        got = out.cpu().numpy().view(np.uint16)
        assert (got == want.view(np.uint16)[:, k0:k0 + 16].T).all(), k0


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("quant", [0, 8, 4])
@pytest.mark.parametrize("M", [1, 16, 40, 64, 200])
def test_linear_stream(dtype, quant, M, monkeypatch):
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "stream")
    ffb_lib.reload_env()
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


# ---- persistent quantised decode kernel (csrc/linear_qgemv.cu): the default for int8 / int4 blobs at <= 16 tokens ----

def _one_hot_columns(m, bdev, K, N, k0s):
    torch = _torch()
    from flexflow_b200 import ops
    got = {}
    for k0 in k0s:
        x = torch.zeros(16, K, dtype=torch.float16, device="cuda")
        x[torch.arange(16), k0 + torch.arange(16)] = 1.0
        out = torch.empty(16, N, dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, out, bdev, None, K, N, 16)
        got[k0] = out.cpu().numpy().view(np.uint16)
    return got


@pytest.mark.parametrize("quant", [8, 4])
@pytest.mark.parametrize("K,N", [(1024, 96), (2048, 160), (4096, 4768)])
def test_linear_qgemv_dequantises_bit_exactly(quant, K, N, monkeypatch):
    """One-hot activations: row m of the output IS column k0+m of the dequantised matrix (fp32 accumulation of one
    non-zero term), so the kernel's on-chip dequantisation is compared bit for bit with the oracle's.  N = 4768 = 149
    groups gives one CTA a second tile (ring wrap-around, coefficient double buffer across a tile boundary)."""
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "qgemv")
    ffb_lib.reload_env()
    rng = np.random.default_rng(K + quant)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant)
    want = O.dequant(blob, quant, K, N).view(np.uint16)
    m = ops.LinearMeta(ops.FFHandler("cuda:0"), K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4,
                       max_tokens_per_batch=16)
    k0s = list(range(0, K, 16)) if K <= 2048 else list(range(0, K, 16))[::5] + [K - 16]
    got = _one_hot_columns(m, torch.from_numpy(blob).cuda(), K, N, k0s)
    for k0 in k0s:
        assert (got[k0] == want[:, k0:k0 + 16].T).all(), k0


@pytest.mark.parametrize("quant", [8, 4])
@pytest.mark.parametrize("M", [1, 7, 16])
@pytest.mark.parametrize("K,N", [(2048, 4128), (4096, 12288)])
def test_linear_qgemv_matches_oracle(quant, M, K, N, monkeypatch):
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "qgemv")
    ffb_lib.reload_env()
    rng = np.random.default_rng(M + K)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    bias = O.round_to_dt(rng.standard_normal(N).astype(np.float32), "half")
    blob = O.quantize(w, quant)
    wq = O.dequant(blob, quant, K, N).view(np.float16).astype(np.float32)
    m = ops.LinearMeta(ops.FFHandler("cuda:0"), K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4,
                       max_tokens_per_batch=16)
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).cuda().half(), out, torch.from_numpy(blob).cuda(),
                                        torch.from_numpy(bias).cuda().half(), K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, wq, bias, "half")
    got = out.float().cpu().numpy()
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())
    # same call again: the fixed-order reduction makes the result reproducible bit for bit
    out2 = torch.empty_like(out)
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).cuda().half(), out2, torch.from_numpy(blob).cuda(),
                                        torch.from_numpy(bias).cuda().half(), K, N, M)
    assert torch.equal(out, out2)


@pytest.mark.parametrize("quant", [8, 4])
def test_linear_qgemv_zero_scale_takes_the_elementwise_rule(quant, monkeypatch):
    """A scale of 0 makes every weight of its column inf / NaN in the reference (q / 0 + offset); the stage that holds
    it falls back to the element-wise rule, the other groups and the group's other stages stay on the pair rule."""
    torch = _torch()
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "qgemv")
    ffb_lib.reload_env()
    K, N = 2048, 96
    rng = np.random.default_rng(3)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant).copy()
    nvals = K * N * quant // 8
    scales = blob[nvals + 2 * (K * N // 32):].view(np.float16)
    scales[1 * K + 5] = 0.0   # group 1, column 5 (stage 0)
    want = O.dequant(blob, quant, K, N).view(np.uint16)
    m = ops.LinearMeta(ops.FFHandler("cuda:0"), K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4,
                       max_tokens_per_batch=16)
    k0s = [0, 16, 1024, 2032]
    got = _one_hot_columns(m, torch.from_numpy(blob).cuda(), K, N, k0s)
    for k0 in k0s:
        ok = np.r_[0:32, 64:96]
        assert (got[k0][:, ok] == want[ok, k0:k0 + 16].T).all(), k0
        g1 = got[k0][:, 32:64].view(np.float16)
        assert not np.isfinite(g1).any()  # inf, or 0 * inf = NaN, in every dot product of the group


# ---- the reference's OWN decompress kernels (oracle/_ref/libref_decompress.so, compiled from
# /root/reference/src/ops/kernels/decompress_kernels.cu by oracle/Makefile) run on the same GPU ----

def _ref_dequant(blob_dev, bits, in_dim, out_dim):
    torch = _torch()
    ref = O.ref_lib()
    out = torch.empty(out_dim, in_dim, dtype=torch.float16, device="cuda")
    fn = ref.ref_decompress_int4_half if bits == 4 else ref.ref_decompress_int8_half
    rc = fn(C.c_void_p(blob_dev.data_ptr()), C.c_void_p(out.data_ptr()), in_dim, out_dim,
            C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    return out


needs_ref = pytest.mark.skipif(O.ref_lib() is None, reason="oracle/_ref/libref_decompress.so not built (make -C oracle ref)")


@needs_ref
@pytest.mark.parametrize("bits", [4, 8])
@pytest.mark.parametrize("shape", [(12288, 4096), (4096, 4096), (11008, 4096), (4096, 11008)])
def test_dequant_bit_exact_against_the_reference_kernel(bits, shape):
    """the four LLaMA-2-7B weight shapes, random values and random fp16 scales / offsets of both
    signs and all magnitudes a compress() blob can hold; ffb_dequant_* == the reference kernel, every bit."""
    torch = _torch()
    from flexflow_b200 import ops
    out_dim, in_dim = shape
    g = torch.Generator(device="cuda").manual_seed(bits * 1000 + out_dim)
    n = in_dim * out_dim
    vals = torch.randint(0, 256, (n * bits // 8,), generator=g, device="cuda", dtype=torch.uint8)
    sc = (torch.randn(n // 32, generator=g, device="cuda") * torch.exp(6 * torch.randn(n // 32, generator=g, device="cuda"))).half()
    sc[sc == 0] = 1.0  # zero scales have their own test below
    of = (torch.randn(n // 32, generator=g, device="cuda") * 0.1).half()
    of[::13] = -0.0  # with q = 0 and a negative scale the reference gives -0 + -0 = -0 (sign of the zero quotient)
    of[5::13] = 0.0
    vals[::7] = 0
    blob = torch.cat([vals, of.view(torch.uint8), sc.view(torch.uint8)])
    want = _ref_dequant(blob, bits, in_dim, out_dim)
    got = ops.decompress_weights(blob, in_dim, out_dim, B.DT_INT4 if bits == 4 else B.DT_INT8)
    torch.cuda.synchronize()
    assert torch.equal(got.view(torch.int16), want.view(torch.int16))


@needs_ref
@pytest.mark.parametrize("bits", [4, 8])
def test_oracle_and_kernels_agree_with_the_reference_kernel_on_edge_scales(bits):
    """Pins the CPU oracle itself: oracle == reference kernel == ffb_dequant on a blob whose scales sweep every finite
    half (zero, subnormals, both signs, inf) and whose offsets include inf."""
    torch = _torch()
    from flexflow_b200 import ops
    in_dim, out_dim = 65536 // 8, 32 * 8  # 65536 groups-columns: one per half bit pattern
    n = in_dim * out_dim
    rng = np.random.default_rng(bits)
    vals = rng.integers(0, 256, n * bits // 8, dtype=np.uint8)
    sc = np.arange(65536, dtype=np.uint16)
    sc = sc[~np.isnan(sc.view(np.float16))]  # NaN payload propagation is not part of the contract
    sc = np.resize(sc, n // 32)
    of = (rng.standard_normal(n // 32) * 0.1).astype(np.float16)
    of[::997] = np.inf
    of[1::5] = -0.0
    of[2::5] = 0.0
    vals[::3] = 0
    blob = np.concatenate([vals, of.view(np.uint8), sc.view(np.uint8)])
    bdev = torch.from_numpy(blob).cuda()
    want = _ref_dequant(bdev, bits, in_dim, out_dim).cpu().numpy().view(np.uint16)
    got = ops.decompress_weights(bdev, in_dim, out_dim, B.DT_INT4 if bits == 4 else B.DT_INT8).cpu().numpy().view(np.uint16)
    orc = O.dequant(blob, bits, in_dim, out_dim).view(np.uint16)
    nan_w = np.isnan(want.view(np.float16))
    for name, x in (("kernel", got), ("oracle", orc)):
        assert (np.isnan(x.view(np.float16)) == nan_w).all(), name
        assert (x[~nan_w] == want[~nan_w]).all(), name


# ---- the reference Linear's own GEMM call (cublasGemmEx, fp16 accumulation) restated in oracle/_ref/libref_linear.so ----
needs_ref_linear = pytest.mark.skipif(O.ref_linear_lib() is None, reason="oracle/_ref/libref_linear.so not built")


@needs_ref_linear
@pytest.mark.parametrize("M", [1, 16, 64, 200, 512])
@pytest.mark.parametrize("K,N", [(4096, 12288), (4096, 4096)])
def test_linear_against_the_reference_cublas_call(M, K, N):
    """Every dispatch range of ffb_linear_forward (GEMV, tcgen05 1-SM and SM-pair kernels) against the call the
    reference makes.  The reference accumulates in fp16 (compute type = DT), the product in fp32, so the product must
    sit closer to the fp32 oracle than the reference does, and the two must agree within the reference's own error."""
    torch = _torch()
    from flexflow_b200 import ops
    ref = O.ref_linear_lib()
    g = torch.Generator(device="cuda").manual_seed(M + N)
    x = torch.randn(M, K, generator=g, device="cuda").half()
    w = (torch.randn(N, K, generator=g, device="cuda") * 0.02).half()
    m = ops.LinearMeta(ops.FFHandler("cuda:0"), K, N, max_tokens_per_batch=1024)
    got = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, x, got, w, None, K, N, M)
    want = torch.empty(M, N, dtype=torch.float16, device="cuda")
    assert ref.ref_linear_forward_half(C.c_void_p(x.data_ptr()), C.c_void_p(w.data_ptr()), C.c_void_p(want.data_ptr()), K, N, M,
                                       C.c_void_p(torch.cuda.current_stream().cuda_stream)) == 0
    torch.cuda.synchronize()
    exact = x.double() @ w.double().t()
    scale = exact.abs().max().item()
    err_ours = (got.double() - exact).abs().max().item() / scale
    err_ref = (want.double() - exact).abs().max().item() / scale
    assert err_ours < 2e-3, err_ours
    assert err_ours <= err_ref + 1e-4, (err_ours, err_ref)
    assert (got.double() - want.double()).abs().max().item() / scale < 1e-2 + 2 * err_ref


@pytest.mark.parametrize("quant", [0, 8, 4])
@pytest.mark.parametrize("M", [16, 64, 512])
@pytest.mark.parametrize("K,N", [(4096, 11008), (11008, 4096)])
def test_linear_mlp_projection_shapes(K, N, M, quant):
    """gate / up (4096 -> 11008) and down (11008 -> 4096) of the LLaMA-2-7B MLP go through the same
    ffb_linear_forward (the reference runs them through the same Kernels::Linear::forward_kernel): every dispatch range,
    dense and quantised, against a torch fp32 product of the same (dequantised) 16-bit weights."""
    torch = _torch()
    from flexflow_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(K + N + M + quant)
    x = torch.randn(M, K, generator=g, device="cuda", dtype=torch.float16)
    w = torch.randn(N, K, generator=g, device="cuda") * 0.02
    h = ops.FFHandler("cuda:0")
    if quant == 0:
        wdev, qt, wref = w.half(), B.DT_NONE, w.half().float()
    else:
        blob = O.quantize(w.cpu().numpy(), quant)
        wdev, qt = torch.from_numpy(blob).cuda(), (B.DT_INT8 if quant == 8 else B.DT_INT4)
        wref = ops.decompress_weights(wdev, K, N, qt).float()
    m = ops.LinearMeta(h, K, N, torch.float16, qt, 512)
    out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, x, out, wdev, None, K, N, M)
    torch.cuda.synchronize()
    want = x.float() @ wref.t()
    assert torch.isfinite(out.float()).all()
    assert float((out.float() - want).abs().max() / want.abs().max()) < 4e-3
