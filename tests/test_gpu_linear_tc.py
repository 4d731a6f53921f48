"""GPU parity: the tcgen05/TMA projection GEMM (M >= 17 for dense weights) vs the CPU oracle.  Kept in its own file so the
driver script can run it under a process-level timeout (a wrong mbarrier phase hangs rather than fails)."""
import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as ffb_lib  # noqa: E402
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("M,K,N", [(17, 1024, 1536), (40, 512, 8192), (65, 1024, 1536), (128, 1024, 1536), (200, 512, 256), (512, 2048, 1024), (1000, 1024, 128)])
def test_linear_tc_fp(dtype, M, K, N, monkeypatch):
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    ffb_lib.reload_env()
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    rng = np.random.default_rng(31)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), dtype)
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), dtype)
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), dtype)
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=tdt, max_tokens_per_batch=1024)
    out = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).to("cuda", tdt), out,
                                        torch.from_numpy(w).to("cuda", tdt), torch.from_numpy(b).to("cuda", tdt), K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, w, b, dtype)
    got = out.float().cpu().numpy()
    assert np.isfinite(got).all()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert np.abs(got - want).max() <= tol * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("bn", ["128", "256", "168", "80", "auto"])
@pytest.mark.parametrize("M,K,N", [(129, 512, 256), (300, 1024, 1024), (512, 4096, 512), (1024, 1024, 768), (512, 512, 1248)])
def test_linear_tc2_pair_tiles(M, K, N, bn, monkeypatch):
    """The SM-pair (cta_group::2) kernel at both tile widths, ragged M (TMA zero-fill + row guard)."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    ffb_lib.reload_env()
    if bn != "auto":
        monkeypatch.setenv("FFB_TC2_BN", bn)  # widths that do not divide N exercise the partial last column tile
        ffb_lib.reload_env()
    rng = np.random.default_rng(33)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), "half")
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=torch.float16, max_tokens_per_batch=1024)
    out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).cuda().half(), out, torch.from_numpy(w).cuda().half(),
                                        None, K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, w, None, "half")
    got = out.float().cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("split", ["2", "auto"])
@pytest.mark.parametrize("M,K,N,bias", [(512, 4096, 1024, True), (300, 2048, 768, False), (200, 1024, 1248, True),
                                        (512, 4096, 4096, False)])
def test_linear_tc2_split_k(M, K, N, bias, split, monkeypatch):
    """Split-K SM-pair GEMM (the second k-slice parks its fp32 partial in the L2 workspace, the tile's owner bulk-copies
    it into its idle operand ring and adds it), called three times on one handle: the arrival counters must come back
    to zero and the result must not depend on timing."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    ffb_lib.reload_env()
    if split != "auto":
        monkeypatch.setenv("FFB_TC2_SPLIT", split)
        ffb_lib.reload_env()
        monkeypatch.setenv("FFB_TC2_BN", "256" if N >= 1024 else "128")
        ffb_lib.reload_env()
    rng = np.random.default_rng(35)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), "half")
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), "half") if bias else None
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=torch.float16, max_tokens_per_batch=1024)
    xd, wd = torch.from_numpy(x).cuda().half(), torch.from_numpy(w).cuda().half()
    bd = torch.from_numpy(b).cuda().half() if bias else None
    outs = []
    for _ in range(3):
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, xd, out, wd, bd, K, N, M)
        outs.append(out)
    torch.cuda.synchronize()
    want = O.linear(x, w, b, "half")
    got = outs[0].float().cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("quant", [8, 4])
def test_linear_tc_quantised_matches_stream(quant, monkeypatch):
    """Large-M quantised path (dequant scratch + tcgen05 GEMM) against the oracle on the dequantised weights."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    ffb_lib.reload_env()
    M, K, N = 300, 1024, 512
    rng = np.random.default_rng(32)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant)
    wq = O.dequant(blob, quant, K, N).view(np.float16).astype(np.float32)
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4, max_tokens_per_batch=512)
    out = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, torch.from_numpy(x).cuda().half(), out, torch.from_numpy(blob).cuda(),
                                        None, K, N, M)
    torch.cuda.synchronize()
    want = O.linear(x, wq, None, "half")
    assert np.abs(out.float().cpu().numpy() - want).max() <= 2e-3 * max(1.0, np.abs(want).max())


@pytest.mark.parametrize("quant", [8, 4])
@pytest.mark.parametrize("M", [65, 300])
def test_linear_tcq_dequantises_bit_exactly_in_shared_memory(quant, M, monkeypatch):
    """linear_tcq.cu (M > 64, quantised weights dequantised in shared memory ahead of the tcgen05 MMA): with one-hot
    activations row m of the output IS column k0 + m of the dequantised matrix (a single fp32 term), so the converter
    warps' output is compared bit for bit with the oracle's dequantisation - values of every code, scales of both
    signs and an infinite one (the element-wise fallback), every k-block, ragged last M tile; plus a
    bias."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_TCQ", "1")  # off by default: slower than dequantise-to-scratch + GEMM (profiles/README.md)
    ffb_lib.reload_env()
    K, N = 512, 256
    rng = np.random.default_rng(quant * 100 + M)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant).copy()
    nv = N * K // (2 if quant == 4 else 1)
    meta = blob[nv:].view(np.uint16)
    ng = N * K // 32
    sc = meta[ng:]
    # an infinite and a negative scale (both take the element-wise fallback and stay finite).  A zero scale is left out:
    # its column dequantises to inf / NaN and 0 x inf would poison every output of those features through the MMA.
    sc[77], sc[300] = 0x7C00, (sc[300] | 0x8000)
    want = O.dequant(blob, quant, K, N).view(np.float16)  # [N][K]
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4, max_tokens_per_batch=512)
    bdev = torch.from_numpy(blob).cuda()
    bias = torch.from_numpy(rng.standard_normal(N).astype(np.float16)).cuda()
    for k0 in range(0, K, M):
        n = min(M, K - k0)
        x = torch.zeros(M, K, dtype=torch.float16, device="cuda")
        x[torch.arange(n), k0 + torch.arange(n)] = 1.0
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, out, bdev, None, K, N, M)
        got = out[:n].cpu().numpy()
        exp = want[:, k0:k0 + n].T
        fin = np.isfinite(exp.astype(np.float32))
        bad = (got.view(np.uint16) != exp.view(np.uint16)) & fin
        assert not bad.any(), (k0, int(bad.sum()), np.argwhere(bad)[:8].tolist(), got[bad][:4], exp[bad][:4])
        assert not np.isfinite(got.astype(np.float32)[~fin]).any()
        if n < M:
            assert (out[n:].float() == 0).all()
    out_b = torch.empty(M, N, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, x, out_b, bdev, bias, K, N, M)
    torch.cuda.synchronize()
    ok = torch.isfinite(out.float()) & torch.isfinite(out_b.float())
    assert torch.equal(out_b[ok], (out.float() + bias.float()[None, :]).half()[ok])


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("M,K,N,bias", [(512, 4096, 12288, False), (512, 4096, 4096, True), (256, 2048, 4096, True),
                                        (300, 1024, 1248, True), (1024, 1024, 2048, False), (129, 4096, 512, False)])
def test_linear_sk_stream_k(M, K, N, bias, dtype, monkeypatch):
    """Stream-K schedule of the SM-pair GEMM (linear_sk.cu, forced with FFB_SK=1): pieces that start / end inside a
    tile, tiles cut into two and into many pieces (k-block ranges shorter than K/64), ragged M and N; called three
    times on one handle: the flags must come back to zero and the result must not depend on timing."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    monkeypatch.setenv("FFB_SK", "1")
    ffb_lib.reload_env()
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    rng = np.random.default_rng(41)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), dtype)
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), dtype)
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), dtype) if bias else None
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=tdt, max_tokens_per_batch=1024)
    xd, wd = torch.from_numpy(x).to("cuda", tdt), torch.from_numpy(w).to("cuda", tdt)
    bd = torch.from_numpy(b).to("cuda", tdt) if bias else None
    outs = []
    for _ in range(3):
        out = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, xd, out, wd, bd, K, N, M)
        outs.append(out)
    torch.cuda.synchronize()
    want = torch.from_numpy(x).cuda() @ torch.from_numpy(w).cuda().T  # fp32 on the rounded inputs
    if bias:
        want = want + torch.from_numpy(b).cuda()[None, :]
    got = outs[0].float()
    assert torch.isfinite(got).all()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert (got - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("M,K,N,bias,bn", [(512, 4096, 12288, False, None), (300, 1024, 1024, True, None),
                                           (1024, 1024, 2048, True, None), (1024, 512, 20480, False, 512),
                                           (129, 2048, 512, True, None), (512, 1024, 11008, True, None),
                                           (512, 512, 12288, False, 400), (260, 1024, 4128, True, 464)])
def test_linear_w512_tiles(M, K, N, bias, bn, dtype, monkeypatch):
    """One wave of wide tiles per SM pair (linear_w512.cu, forced with FFB_W512=1): 256 x BN tiles, 272 <= BN <= 512 chosen
    so that all tiles fit the 74 pairs (336 for 512 x 12288, 304 for 512 x 11008), two UMMAs of unequal width per k-step
    against one A tile, single-buffered accumulator; ragged M and ragged last column block (N = 4128), forced widths
    with several tiles per pair (160 tiles of 512 on 74 pairs: the accumulator hand-over between tiles), the
    programmatic-dependent-launch prologue on back-to-back calls."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    monkeypatch.setenv("FFB_W512", "1")
    if bn is not None:
        monkeypatch.setenv("FFB_W512_BN", str(bn))
    ffb_lib.reload_env()
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    g = torch.Generator(device="cuda").manual_seed(43)
    x = torch.randn(M, K, generator=g, device="cuda").to(tdt)
    w = (torch.randn(N, K, generator=g, device="cuda") * 0.05).to(tdt)
    b = torch.randn(N, generator=g, device="cuda").to(tdt) if bias else None
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=tdt, max_tokens_per_batch=1024)
    outs = []
    for _ in range(3):
        out = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, out, w, b, K, N, M)
        outs.append(out)
    torch.cuda.synchronize()
    want = x.float() @ w.float().T
    if bias:
        want = want + b.float()[None, :]
    got = outs[0].float()
    assert torch.isfinite(got).all()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert (got - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])


@pytest.mark.parametrize("dtype", ["half", "bf16"])
@pytest.mark.parametrize("M,K,N,bias,bn,split", [(64, 4096, 4096, False, None, None), (17, 1024, 12288, True, None, None),
                                                 (128, 2048, 4128, True, None, None), (100, 11008, 4096, False, None, None),
                                                 (33, 512, 256, True, None, None), (48, 4096, 32000, False, None, None),
                                                 (64, 4096, 12288, False, None, None), (128, 4096, 11008, True, None, None),
                                                 (128, 1024, 1024, True, 128, 8), (40, 4096, 2048, True, 128, 2),
                                                 (96, 1024, 4096, False, 128, 1), (72, 2048, 4128, True, 192, 4),
                                                 (128, 1024, 2048, True, 256, 8), (24, 512, 1280, False, 256, 1),
                                                 (512, 4096, 1536, False, 128, 3), (256, 4096, 4096, True, None, None),
                                                 (300, 1024, 1024, True, None, None), (200, 2048, 256, False, None, None),
                                                 (100, 1536, 1024, True, 128, 3), (384, 2048, 2080, True, 128, 3)])
def test_linear_tcc_cluster_split_k(M, K, N, bias, bn, split, dtype, monkeypatch):
    """17..128 tokens: one tile of 128 / 192 / 256 features per thread-block cluster, the reduction dimension split over
    its 2 / 4 / 8 CTAs and summed through distributed shared memory in rank order (linear_tcc.cu): o_proj (32 tiles x 4),
    qkv_proj (64 tiles of 192 x 2), K = 11008 (172 k-blocks over 4 CTAs: uneven ranges), tiny layers (2 tiles x 8), no
    split (N = 32000), ragged N (4128), token quarters without rows (M = 17, 24, 33, 40); more than 128 tokens on narrow
    layers (o_proj at 256 tokens: 2 token blocks x 32 tiles x 2; forced: 4 x 12 tiles x clusters of 3 with columns 48 / 48 / 32;
    a ragged last token block, M = 300); three calls on one handle give the same bits.  Against fp32 torch on the rounded inputs and against linear_tc's one-CTA-per-tile result."""
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "tc")
    if split is not None:
        monkeypatch.setenv("FFB_TCC_BN", str(bn))
        monkeypatch.setenv("FFB_TCC_SPLIT", str(split))
    ffb_lib.reload_env()
    tdt = torch.float16 if dtype == "half" else torch.bfloat16
    rng = np.random.default_rng(47)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), dtype)
    w = O.round_to_dt((rng.standard_normal((N, K)) * 0.05).astype(np.float32), dtype)
    b = O.round_to_dt(rng.standard_normal(N).astype(np.float32), dtype) if bias else None
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, K, N, dtype=tdt, max_tokens_per_batch=512)
    xd, wd = torch.from_numpy(x).to("cuda", tdt), torch.from_numpy(w).to("cuda", tdt)
    bd = torch.from_numpy(b).to("cuda", tdt) if bias else None
    outs = []
    for _ in range(3):
        out = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, xd, out, wd, bd, K, N, M)
        outs.append(out)
    torch.cuda.synchronize()
    want = torch.from_numpy(x).cuda() @ torch.from_numpy(w).cuda().T  # fp32 on the rounded inputs
    if bias:
        want = want + torch.from_numpy(b).cuda()[None, :]
    got = outs[0].float()
    assert torch.isfinite(got).all()
    tol = 2e-3 if dtype == "half" else 1.2e-2
    assert (got - want).abs().max().item() <= tol * max(1.0, want.abs().max().item())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    # the one-CTA-per-tile kernel (no k split; the SM-pair kernel above 128 tokens) on the same inputs: equal up to the rounding of differently ordered fp32 sums
    monkeypatch.setenv("FFB_TCC", "0")
    ffb_lib.reload_env()
    ref = torch.full((M, N), float("nan"), dtype=tdt, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, xd, ref, wd, bd, K, N, M)
    torch.cuda.synchronize()
    assert (ref.float() - got).abs().max().item() <= tol * max(1.0, want.abs().max().item())
    monkeypatch.delenv("FFB_TCC")
    ffb_lib.reload_env()
