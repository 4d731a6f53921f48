"""GPU: weight files -> blob -> kernels.  A matrix quantised by ffb_quantize_weights, written as the
reference's three files, read back by ffb_load_from_quantized_file and dequantised by ffb_dequant_int4 / _int8 on the
GPU equals the oracle's dequantisation of the directly built blob bit for bit, and the quantised Linear forward fed
with the loaded blob equals the dense product of the dequantised matrix; the dense qkv_proj / o_proj shards read from
HF-export files drive ffb_linear_forward to the result of the unsharded matrices."""
import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import file_loader as FL
from oracle import oracle as O

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("bits", [8, 4])
def test_quantise_save_load_dequant_round_trip(bits, tmp_path):
    import torch
    from flexflow_b200 import ops
    rng = np.random.default_rng(bits)
    in_dim, out_dim = 1024, 512
    w = (rng.standard_normal((out_dim, in_dim)) * 0.03).astype(np.float32)
    qt = B.DT_INT4 if bits == 4 else B.DT_INT8
    blob = FL.compress(w, bits)
    FL.save_quantized(str(tmp_path / "w"), blob, in_dim, out_dim, bits)
    loaded = FL.FileDataLoader(str(tmp_path), 1, 1, in_dim, 1).load_from_quantized_file("w", in_dim, out_dim, qt)
    dq = ops.decompress_weights(torch.from_numpy(loaded).cuda(), in_dim, out_dim, qt)
    torch.cuda.synchronize()
    want = O.dequant(blob, bits, in_dim, out_dim)
    assert np.array_equal(dq.cpu().numpy().view(np.uint16), want)
    step = (w.reshape(-1, 32, in_dim).max(1) - w.reshape(-1, 32, in_dim).min(1)) / (2 ** bits - 1)
    err = np.abs(O.from_bits(want) - w).reshape(-1, 32, in_dim)
    assert (err <= step[:, None, :] * 0.51 + 1e-3 * np.abs(w).max()).all()
    # the quantised Linear forward on the loaded blob (decode batch: dequantisation inside the weight stream)
    h = ops.FFHandler("cuda:0")
    m = ops.LinearMeta(h, in_dim, out_dim, torch.float16, qt, 16)
    x = torch.from_numpy((rng.standard_normal((16, in_dim))).astype(np.float16)).cuda()
    y = torch.empty(16, out_dim, dtype=torch.float16, device="cuda")
    ops.Linear.inference_kernel_wrapper(m, None, x, y, torch.from_numpy(loaded).cuda(), None, in_dim, out_dim, 16)
    torch.cuda.synchronize()
    ref = x.float() @ dq.float().t()
    assert float((y.float() - ref).abs().max() / ref.abs().max()) < 5e-3


def test_dense_shards_from_files_reproduce_the_unsharded_projection(tmp_path):
    import torch
    from flexflow_b200 import ops
    rng = np.random.default_rng(1)
    nh, nkv, dh, hidden, tp = 8, 2, 64, 512, 2
    wq = (rng.standard_normal((nh * dh, hidden)) * 0.05).astype(np.float16)
    wk = (rng.standard_normal((nkv * dh, hidden)) * 0.05).astype(np.float16)
    wv = (rng.standard_normal((nkv * dh, hidden)) * 0.05).astype(np.float16)
    wo = (rng.standard_normal((hidden, nh * dh)) * 0.05).astype(np.float16)
    for s, a in ((".q_proj.weight", wq), (".k_proj.weight", wk), (".v_proj.weight", wv), (".o_proj.weight", wo)):
        a.tofile(str(tmp_path / ("layers.0.self_attn" + s)))
    ld = FL.FileDataLoader(str(tmp_path), nh, nkv, hidden, dh, tp)
    h = ops.FFHandler("cuda:0")
    x = torch.from_numpy(rng.standard_normal((24, hidden)).astype(np.float16)).cuda()
    a = torch.from_numpy(rng.standard_normal((24, nh * dh)).astype(np.float16)).cuda()
    H_l = nh // tp * dh
    o_sum = torch.zeros(24, hidden, device="cuda")
    for r in range(tp):
        wqkv = torch.from_numpy(ld.load_attention_weights_to_dense("layers.0.self_attn", False, r).view(np.float16)).cuda()
        m = ops.LinearMeta(h, hidden, 3 * H_l, torch.float16, B.DT_NONE, 32)
        y = torch.empty(24, 3 * H_l, dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(m, None, x, y, wqkv, None, hidden, 3 * H_l, 24)
        # shard r: [Q heads of r | their kv heads, replicated | likewise V]
        heads = range(r * nh // tp, (r + 1) * nh // tp)
        tq = torch.from_numpy(wq).cuda().float()
        tk = torch.from_numpy(wk).cuda().float()
        want_q = x.float() @ torch.cat([tq[hh * dh:(hh + 1) * dh] for hh in heads]).t()
        want_k = x.float() @ torch.cat([tk[(hh // (nh // nkv)) * dh:(hh // (nh // nkv) + 1) * dh] for hh in heads]).t()
        torch.cuda.synchronize()
        assert float((y[:, :H_l].float() - want_q).abs().max()) < 2e-2
        assert float((y[:, H_l:2 * H_l].float() - want_k).abs().max()) < 2e-2
        wo_r = torch.from_numpy(ld.load_attention_weights_to_dense("layers.0.self_attn", True, r).view(np.float16)).cuda()
        mo = ops.LinearMeta(h, H_l, hidden, torch.float16, B.DT_NONE, 32)
        yo = torch.empty(24, hidden, dtype=torch.float16, device="cuda")
        ops.Linear.inference_kernel_wrapper(mo, None, a[:, r * H_l:(r + 1) * H_l].contiguous(), yo, wo_r, None, H_l, hidden, 24)
        o_sum += yo.float()
    torch.cuda.synchronize()
    want_o = a.float() @ torch.from_numpy(wo).cuda().float().t()
    assert float((o_sum - want_o).abs().max() / want_o.abs().max()) < 1e-2
