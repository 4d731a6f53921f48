"""GPU parity at BASELINE configs[3]'s FULL size, through the whole operator chain: two LLaMA-2-7B attention blocks
(qkv_proj -> TreeIncMultiHeadSelfAttention -> o_proj, 4096 hidden, 32 heads x 128, 8 requests x 64 tree tokens over
1024 committed tokens) run by FFModel.inference - eagerly (programmatic dependent launches between the kernels) and as
a CUDA-graph replay - against an independent recomputation: fp32 torch matmuls for the projections (rounded to half
where the operators round) and the CPU oracle for the attention.  Tolerance 1e-2 relative (north_star).  The
per-operator tests run small shapes; this one would catch an ordering bug between the chained kernels (weights
prefetched ahead of griddepcontrol.wait, early triggers) that only shows at sizes where the kernels overlap."""
import numpy as np
import pytest

import ffb_testlib as TL
from flexflow_b200 import batch_config as B

pytestmark = pytest.mark.gpu

HIDDEN, NQ, DH, NREQ, NTREE, COMMITTED, LAYERS = 4096, 32, 128, 8, 64, 1024, 2


def test_tree_verify_chain_7b_shapes_eager_and_graph():
    import torch
    from flexflow_b200 import ops
    from flexflow_b200.model import FFConfig, FFModel
    from oracle import oracle as O
    dev = "cuda:0"
    T = NREQ * NTREE
    rng = np.random.default_rng(2024)
    g = torch.Generator(device=dev).manual_seed(99)
    cfg = FFConfig(data_type=torch.float16, max_requests_per_batch=NREQ, max_tokens_per_batch=T,
                   max_sequence_length=COMMITTED + 16, max_spec_tree_token_num=NTREE, device=dev)
    model = FFModel(cfg, None)
    t_in = model.create_tensor(HIDDEN)
    t = t_in
    rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=10000.0)
    for i in range(LAYERS):
        qkv = model.dense(t, 3 * HIDDEN, name="l%d.qkv_proj" % i)
        a = model.inc_multiquery_self_attention_verify(qkv, HIDDEN, NQ, NQ, kdim=DH, rotary_embedding_meta=rope,
                                                       name="l%d.attn" % i)
        t = model.dense(a, HIDDEN, name="l%d.o_proj" % i)
    model.compile_inference()
    lins = [l for l in model.layers if l.op_type == "linear"]
    attns = [l for l in model.layers if l.op_type == "tree_attn"]
    for l in lins:
        std = (1.0 / HIDDEN) ** 0.5 * (6.0 if l.name.endswith("o_proj") else 1.0)
        l.weight = (torch.randn(l.output.dim, l.inputs[0].dim, generator=g, device=dev) * std).half()
    # committed K/V rows: the same values in the operator's cache and in the oracle's
    orcs = []
    for l in attns:
        orc = O.OracleAttn(B.TREE_VERIFY_MODE, "half", NQ, DH, NREQ, COMMITTED + 16, B.MAX_SPEC_TREE_TOKEN_NUM, T,
                           apply_rope=True, theta=10000.0)
        for s in range(NREQ):
            k = TL.rand_dt(rng, (COMMITTED, NQ, DH), "half")
            v = TL.rand_dt(rng, (COMMITTED, NQ, DH), "half")
            l.meta.write_kv(s, 0, torch.from_numpy(k).to(dev, torch.float16), torch.from_numpy(v).to(dev, torch.float16))
            orc.kcache[s, :COMMITTED] = k.reshape(COMMITTED, NQ * DH)
            orc.vcache[s, :COMMITTED] = v.reshape(COMMITTED, NQ * DH)
        orcs.append(orc)
    bc = TL.make_tree_batch([dict(slot=s, committed_len=COMMITTED, parent=TL.random_tree(rng, NTREE)) for s in range(NREQ)])
    assert bc.num_tokens == T
    x0 = torch.randn(T, HIDDEN, generator=g, device=dev).half()

    # ---- independent recomputation ----
    x = x0
    for i in range(LAYERS):
        qkv_ref = (x.float() @ lins[2 * i].weight.float().T).half()
        attn_ref = orcs[i].tree_forward(bc, qkv_ref.float().cpu().numpy().reshape(T, 3, HIDDEN))
        a = torch.from_numpy(np.ascontiguousarray(attn_ref, dtype=np.float32)).to(dev).half()
        x = (a.float() @ lins[2 * i + 1].weight.float().T).half()
    want = x.float()
    scale = want.abs().max().item()
    assert scale > 0.05  # the chain did not collapse to zero

    # ---- product: eager (twice: the second step re-appends the same tree rows) and graph replay ----
    outs = []
    for _ in range(2):
        outs.append(model.inference(bc, t_in, x0, upload=True).float().clone())
    torch.cuda.synchronize()
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        out_g = model.inference(bc, t_in, x0, upload=False)
    for _ in range(2):
        graph.replay()
    torch.cuda.synchronize()
    outs.append(out_g.float().clone())
    for got in outs:
        assert torch.isfinite(got).all()
        err = (got - want).abs().max().item() / scale
        assert err < 1e-2, err
    # the same launches give the same bits, eager or replayed
    assert torch.equal(outs[0], outs[1]) and torch.equal(outs[0], outs[2])
    for l in model.layers:
        if l.meta is not None and hasattr(l.meta, "close"):
            l.meta.close()
