"""GPU parity of linear_qtc.cu: quantised (int8 / int4) projection for up to 64 tokens - TMA stream of the blob,
dequantisation in shared memory, tcgen05 MMA with the weights as the 128-row operand - through the C ABI, against the
CPU oracle.  Dequantised values are compared BIT FOR BIT (one-hot activations: an output row is one column of the
dequantised matrix); whole products within the fp16 tolerance of the other projection tests."""
import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as ffb_lib
from oracle import oracle as O

pytestmark = pytest.mark.gpu


def _meta(K, N, quant, max_tokens=64):
    from flexflow_b200 import ops
    return ops.LinearMeta(ops.FFHandler("cuda:0"), K, N, quantization_type=B.DT_INT8 if quant == 8 else B.DT_INT4,
                          max_tokens_per_batch=max_tokens)


def _one_hot_columns(m, bdev, K, N, k0s, M=16):
    import torch
    from flexflow_b200 import ops
    got = {}
    for k0 in k0s:
        x = torch.zeros(M, K, dtype=torch.float16, device="cuda")
        x[torch.arange(M), k0 + torch.arange(M)] = 1.0
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, out, bdev, None, K, N, M)
        got[k0] = out.cpu().numpy().view(np.uint16)
    return got


@pytest.mark.parametrize("quant", [8, 4])
@pytest.mark.parametrize("K,N,M", [(1024, 2432, 16), (4096, 640, 16), (2048, 1280, 48)])
def test_qtc_dequantises_bit_exactly(quant, K, N, M, monkeypatch):
    """Every k-step of every tile, tiles cut between CTAs at arbitrary k-steps (19 x 8, 5 x 32 and 10 x 16 steps over
    148 CTAs), int4 nibble order, padded token counts (48 -> three 16-column MMA blocks)."""
    import torch
    monkeypatch.setenv("FFB_LINEAR", "qtc")
    ffb_lib.reload_env()
    rng = np.random.default_rng(K + quant + M)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant)
    want = O.dequant(blob, quant, K, N).view(np.uint16)
    m = _meta(K, N, quant)
    k0s = list(range(0, K - M + 1, M))
    if len(k0s) > 48:
        k0s = k0s[::5]
    k0s.append(K - M)
    got = _one_hot_columns(m, torch.from_numpy(blob).cuda(), K, N, k0s, M)
    for k0 in k0s:
        assert (got[k0] == want[:, k0:k0 + M].T).all(), k0


@pytest.mark.parametrize("quant", [8, 4])
def test_qtc_degenerate_scales(quant, monkeypatch):
    """Infinite and negative scales take the element-wise rule for their k-step (bit-exact, finite); a zero scale makes
    its column inf / NaN in the reference, which poisons every product of that FEATURE group only."""
    import torch
    monkeypatch.setenv("FFB_LINEAR", "qtc")
    ffb_lib.reload_env()
    K, N = 1024, 2432
    rng = np.random.default_rng(5)
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    blob = O.quantize(w, quant).copy()
    nvals = K * N * quant // 8
    scales = blob[nvals + 2 * (K * N // 32):].view(np.uint16)
    scales[3 * K + 200] = 0x7C00          # group 3, column 200: +inf
    scales[40 * K + 17] |= 0x8000         # group 40, column 17: negative
    want = O.dequant(blob, quant, K, N).view(np.uint16)
    m = _meta(K, N, quant)
    k0s = [0, 16, 192, 1008]
    got = _one_hot_columns(m, torch.from_numpy(blob).cuda(), K, N, k0s)
    for k0 in k0s:
        assert (got[k0] == want[:, k0:k0 + 16].T).all(), k0
    scales[7 * K + 130] = 0               # group 7 (features 224..255), column 130: zero
    want = O.dequant(blob, quant, K, N).view(np.uint16)
    got = _one_hot_columns(m, torch.from_numpy(blob).cuda(), K, N, [128])[128]
    ok = np.r_[0:224, 256:N]
    assert (got[:, ok] == want[ok, 128:144].T).all()
    assert not np.isfinite(got[:, 224:256].view(np.float16)).any()


@pytest.mark.parametrize("quant", [8, 4])
@pytest.mark.parametrize("M", [1, 7, 16, 40, 64])
@pytest.mark.parametrize("K,N", [(2048, 4224), (4096, 12288)])
def test_qtc_matches_oracle(quant, M, K, N, monkeypatch):
    import torch
    from flexflow_b200 import ops
    monkeypatch.setenv("FFB_LINEAR", "qtc")
    ffb_lib.reload_env()
    rng = np.random.default_rng(M + K)
    x = O.round_to_dt(rng.standard_normal((M, K)).astype(np.float32), "half")
    w = (rng.standard_normal((N, K)) * 0.05).astype(np.float32)
    bias = O.round_to_dt(rng.standard_normal(N).astype(np.float32), "half")
    blob = O.quantize(w, quant)
    wq = O.dequant(blob, quant, K, N).view(np.float16).astype(np.float32)
    m = _meta(K, N, quant)
    xd, bd, biasd = torch.from_numpy(x).cuda().half(), torch.from_numpy(blob).cuda(), torch.from_numpy(bias).cuda().half()
    outs = []
    for _ in range(3):  # the flags of the cut tiles must come back to zero; fixed-order merges: same bits every time
        out = torch.full((M, N), float("nan"), dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, xd, out, bd, biasd, K, N, M)
        outs.append(out)
    torch.cuda.synchronize()
    want = O.linear(x, wq, bias, "half")
    got = outs[0].float().cpu().numpy()
    assert np.isfinite(got).all()
    assert np.abs(got - want).max() <= 2e-3 * max(1.0, np.abs(want).max())
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
