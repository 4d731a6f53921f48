"""GPU: ffb_inc_attn_forward (decode) against the REFERENCE's own kernel, compute_attention_kernel_generation_kernel
(src/ops/inc_multihead_self_attention.cu:372-623), compiled for sm_100a from the reference tree into
oracle/_ref/libref_attn.so (oracle/Makefile target `ref`) and run on the same GPU with the same inputs.  This pins the
CPU oracle's restatement of that kernel as well.  RoPE is off in these cases: the reference kernel reads queries that
compute_qkv_kernel has already rotated in place, which the product keeps private (RoPE itself is checked against the
oracle and the HuggingFace fixtures elsewhere)."""
import ctypes as C

import numpy as np
import pytest

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from oracle import oracle as O

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(O.ref_attn_lib() is None, reason="oracle/_ref/libref_attn.so not built (make -C oracle ref)")


def _reference_decode(p, bc, qkv_dev, lens, max_requests, max_seq):
    """Runs the reference kernel on the product's own cache contents (read back in the reference layout)."""
    import torch
    ref = O.ref_attn_lib()
    assert ref.ref_sizeof_per_request_info() == C.sizeof(B.PerRequestInfo)
    H = p.H
    kc = torch.zeros(max_requests, max_seq, H, dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    for slot, n in lens.items():
        k, v = p.m.read_kv(slot, 0, n + 1)  # committed rows + the token this step appended
        kc[slot, :n + 1] = k.reshape(n + 1, H)
        vc[slot, :n + 1] = v.reshape(n + 1, H)
    ri = torch.from_numpy(np.frombuffer(bytes(bc.requestsInfo), dtype=np.uint8).copy()).cuda()
    T = bc.num_generation_tokens
    out = torch.full((T, H), float("nan"), dtype=torch.float16, device="cuda")
    rc = ref.ref_inc_generation_half(C.c_void_p(qkv_dev.data_ptr()), C.c_void_p(kc.data_ptr()), C.c_void_p(vc.data_ptr()),
                                     C.c_void_p(out.data_ptr()), 1.0 / np.sqrt(p.dh), max_seq, p.dh, H,
                                     C.c_void_p(ri.data_ptr()), p.nq, T,
                                     C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    return out.float().cpu().numpy()


@needs_ref
@pytest.mark.parametrize("dh,nq", [(128, 8), (64, 12)])
@pytest.mark.parametrize("lens", [{0: 5, 1: 130, 2: 777, 3: 2046}, {1: 1, 4: 63, 5: 64, 7: 1500}])
def test_decode_attention_against_the_reference_kernel(dh, nq, lens):
    import torch
    from gpu_harness import Pair, make_qkv
    rng = np.random.default_rng(dh + len(lens))
    max_requests, max_seq = 8, 2048
    p = Pair(B.INC_DECODING_MODE, "half", nq, nq, dh, max_requests, max_seq, 64, apply_rope=False)
    for slot, n in lens.items():
        p.fill_cache(rng, slot, n)
    bc = TL.make_inc_batch([dict(slot=s, first_depth=n, ntok=1) for s, n in lens.items()])
    qkv = make_qkv(rng, len(lens), nq, nq, dh, "half")
    got, orc = p.forward_both(bc, qkv)
    qkv_dev = torch.from_numpy(np.ascontiguousarray(qkv, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    ref = _reference_decode(p, bc, qkv_dev, lens, max_requests, max_seq)
    assert np.isfinite(ref).all()
    # tolerance of north_star: 1e-2 relative in fp16; observed ~1e-3 (the reference sums exp(x - max) with __expf and
    # stores the logits as float in shared memory, the product uses an online softmax in the log2 domain)
    assert TL.rel_err(got, ref) < 1e-2, TL.rel_err(got, ref)
    # and the oracle restates that kernel: closer to it than the product needs to be
    assert TL.rel_err(orc, ref) < 2e-3, TL.rel_err(orc, ref)


@needs_ref
@pytest.mark.parametrize("dh,nq,llama3", [(128, 8, False), (64, 12, False), (128, 4, True)])
def test_rope_kv_append_and_decode_against_the_reference_kernels(dh, nq, llama3):
    """The whole decode chain with RoPE on: the reference's apply_rotary_embedding_hf (inc_...cu:664-738) rotates a
    copy of the projection output in place; its rotated K must be what the product appended to its cache, and the
    reference decode kernel fed with the rotated queries must reproduce the product's output."""
    import torch
    from gpu_harness import Pair, make_qkv
    rng = np.random.default_rng(dh * 3 + nq)
    lens = {0: 17, 2: 511, 3: 1999, 6: 4}
    max_requests, max_seq = 8, 2048
    p = Pair(B.INC_DECODING_MODE, "half", nq, nq, dh, max_requests, max_seq, 64, apply_rope=True, llama3=llama3)
    for slot, n in lens.items():
        p.fill_cache(rng, slot, n)
    bc = TL.make_inc_batch([dict(slot=s, first_depth=n, ntok=1) for s, n in lens.items()])
    T, H = len(lens), nq * dh
    qkv = make_qkv(rng, T, nq, nq, dh, "half")
    got, orc = p.forward_both(bc, qkv)

    ref = O.ref_attn_lib()
    assert ref.ref_sizeof_per_token_info() == C.sizeof(B.PerTokenInfo)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    rot = torch.from_numpy(np.ascontiguousarray(qkv, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    scratch = torch.empty(T * H * 2, dtype=torch.float32, device="cuda")
    ti = torch.from_numpy(np.frombuffer(bytes(bc.tokensInfo), dtype=np.uint8).copy()).cuda()
    rc = ref.ref_apply_rotary_embedding_half(C.c_void_p(rot.data_ptr()), C.c_void_p(scratch.data_ptr()),
                                             C.c_void_p(ti.data_ptr()), 10000.0, int(llama3), 8.0, 1.0, 4.0, 8192, dh, T, H, st)
    assert rc == 0
    torch.cuda.synchronize()
    rot_k = rot[:, 1].float().cpu().numpy().reshape(T, nq, dh)
    # 1. RoPE + KV placement: the row the product appended at abs_depth is the reference's rotated K (and the raw V)
    exact = []
    for t, (slot, n) in enumerate(sorted(lens.items())):
        k, v = p.m.read_kv(slot, n, 1)
        kk = k.float().cpu().numpy()[0]
        exact.append(kk == rot_k[t])
        np.testing.assert_allclose(kk, rot_k[t], rtol=2e-3, atol=2e-3)
        assert np.array_equal(v.float().cpu().numpy()[0].reshape(-1), qkv[t, 2].astype(np.float32))
    assert np.mean(exact) > 0.97, np.mean(exact)  # same expression; sin / cos of the two compilations agree to the last bit almost everywhere
    # 2. attention on the rotated queries
    out_ref = _reference_decode(p, bc, rot, lens, max_requests, max_seq)
    assert TL.rel_err(got, out_ref) < 1e-2, TL.rel_err(got, out_ref)
    assert TL.rel_err(orc, out_ref) < 2e-3, TL.rel_err(orc, out_ref)


@needs_ref
@pytest.mark.parametrize("dh,nq", [(64, 12), (128, 8)])
def test_spec_beam_attention_against_the_reference_kernel(dh, nq):
    """SSM beam step (SpecIncMultiHeadSelfAttention): the reference's own
    compute_spec_inc_attention_kernel_generation_kernel (spec_...cu:36-309) on the product's cache contents, tree shapes
    (1, 1, 3) and (3, 1) as the reference builds them (spec_infer.cc:337-339)."""
    import torch
    from gpu_harness import Pair, make_qkv
    ref = O.ref_attn_lib()
    assert ref.ref_sizeof_beam_per_request_info() == C.sizeof(B.BeamSearchPerRequestInfo)
    assert ref.ref_sizeof_bit_mask() == C.sizeof(B.BitMask)
    rng = np.random.default_rng(dh + 7)
    max_requests, max_seq = 6, 400
    p = Pair(B.BEAM_SEARCH_MODE, "half", nq, nq, dh, max_requests, max_seq, 128, apply_rope=False, quirks=1)
    spec = [dict(slot=0, non_tree=50, widths=[1, 1, 3]), dict(slot=3, non_tree=300, widths=[3, 1])]
    p.fill_cache(rng, 0, 50 + 3)
    p.fill_cache(rng, 3, 300 + 2)
    bc = TL.make_beam_batch(spec)
    T, H = bc.num_tokens, nq * dh
    qkv = make_qkv(rng, T, nq, nq, dh, "half")
    got, orc = p.forward_both(bc, qkv)
    rows = max_seq + B.MAX_SPEC_TREE_TOKEN_NUM  # cache rows in beam mode (inc_...cu:1691-1716)
    kc = torch.zeros(max_requests, rows, H, dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    for r in spec:
        bm = bc.causalMask[r["slot"]]
        n = bm.non_tree_cache_size + bm.tree_size + bm.prompt_size - 1  # totalCacheSize, spec_...cu:93-94
        k, v = p.m.read_kv(r["slot"], 0, n)
        kc[r["slot"], :n] = k.reshape(n, H)
        vc[r["slot"], :n] = v.reshape(n, H)

    def dev(x):
        return torch.from_numpy(np.frombuffer(bytes(x), dtype=np.uint8).copy()).cuda()
    ri, bri, cm, rc_ = dev(bc.requestsInfo), dev(bc.beamRequestsInfo), dev(bc.causalMask), dev(bc.request_completed)
    qkv_dev = torch.from_numpy(np.ascontiguousarray(qkv, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    out = torch.full((T, H), float("nan"), dtype=torch.float16, device="cuda")
    rc = ref.ref_spec_generation_half(C.c_void_p(qkv_dev.data_ptr()), C.c_void_p(kc.data_ptr()), C.c_void_p(vc.data_ptr()),
                                      C.c_void_p(out.data_ptr()), 1.0 / np.sqrt(dh), rows, dh, H, C.c_void_p(ri.data_ptr()),
                                      C.c_void_p(bri.data_ptr()), C.c_void_p(cm.data_ptr()), C.c_void_p(rc_.data_ptr()),
                                      nq, bc.speculative_request_num, C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    out = out.float().cpu().numpy()
    assert np.isfinite(out).all()
    assert TL.rel_err(got, out) < 1e-2, TL.rel_err(got, out)
    assert TL.rel_err(orc, out) < 2e-3, TL.rel_err(orc, out)


def _reference_tree_verify(p, bc, qkv_dev, reqs, max_requests, rows, max_tokens):
    import torch
    ref = O.ref_attn_lib()
    H, T = p.H, bc.num_tokens
    kc = torch.zeros(max_requests, rows, H, dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    max_q = max_total = 0
    for r in reqs:
        n = r["committed_len"] + len(r["parent"])
        k, v = p.m.read_kv(r["slot"], 0, n)  # committed rows + the tree rows this step appended
        kc[r["slot"], :n] = k.reshape(n, H)
        vc[r["slot"], :n] = v.reshape(n, H)
        max_q, max_total = max(max_q, len(r["parent"])), max(max_total, n)

    def dev(x):
        return torch.from_numpy(np.frombuffer(bytes(x), dtype=np.uint8).copy()).cuda()
    ri, cm, rc_ = dev(bc.requestsInfo), dev(bc.causalMask), dev(bc.request_completed)
    out = torch.full((T, H), float("nan"), dtype=torch.float16, device="cuda")
    rc = ref.ref_tree_verify_half(C.c_void_p(qkv_dev.data_ptr()), C.c_void_p(kc.data_ptr()), C.c_void_p(vc.data_ptr()),
                                  C.c_void_p(out.data_ptr()), 1.0 / np.sqrt(p.dh), rows, max_tokens, p.dh, H,
                                  C.c_void_p(ri.data_ptr()), p.nq, len(reqs), C.c_void_p(cm.data_ptr()),
                                  C.c_void_p(rc_.data_ptr()), max_q, max_total,
                                  C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, rc  # -2: the batch would overrun the reference kernel's shared memory (F7) and is not launched
    torch.cuda.synchronize()
    return out.float().cpu().numpy()


@needs_ref
@pytest.mark.parametrize("dh,nq", [(128, 8), (64, 12)])
def test_tree_verify_attention_against_the_reference_kernel(dh, nq):
    """TreeIncMultiHeadSelfAttention: the reference's own compute_attention_kernel_fused_kernel (tree_...cu:35-333) on
    the product's cache contents.  Trees of at most 25 nodes (the reference's `1 << qi` mask shift, F6) and contexts
    short enough for its query-length-sized shared memory (F7: the launcher refuses anything longer); the reference
    carries qk_max across the query tokens of a request, which only moves the softmax's reference point."""
    import torch
    from gpu_harness import Pair, make_qkv
    rng = np.random.default_rng(dh + 11)
    max_requests, max_seq, max_tokens = 8, 1024, 128
    p = Pair(B.TREE_VERIFY_MODE, "half", nq, nq, dh, max_requests, max_seq, max_tokens, apply_rope=False, quirks=1)
    reqs = [dict(slot=0, committed_len=40, parent=TL.tree_from_widths([1, 1, 3])),
            dict(slot=2, committed_len=300, parent=TL.random_tree(rng, 25)),
            dict(slot=5, committed_len=550, parent=TL.random_tree(rng, 25)),
            dict(slot=6, committed_len=7, parent=[-1])]
    for r in reqs:
        p.fill_cache(rng, r["slot"], r["committed_len"])
    bc = TL.make_tree_batch(reqs)
    qkv = make_qkv(rng, bc.num_tokens, nq, nq, dh, "half")
    got, orc = p.forward_both(bc, qkv)
    qkv_dev = torch.from_numpy(np.ascontiguousarray(qkv, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    ref = _reference_tree_verify(p, bc, qkv_dev, reqs, max_requests, max_seq + B.MAX_SPEC_TREE_TOKEN_NUM, max_tokens)
    assert np.isfinite(ref).all()
    assert TL.rel_err(got, ref) < 1e-2, TL.rel_err(got, ref)
    assert TL.rel_err(orc, ref) < 2e-3, TL.rel_err(orc, ref)


@needs_ref
def test_tree_kv_placement_and_commit_against_the_reference_kernels():
    """Two verify steps with a commit in between: the reference's own update_tree_branch_kv_cache_fused
    (tree_...cu:433-478) and commit_tokens_kernel (:335-372) run on a mirror of the cache in the reference layout; after
    each step every cache row of every request equals the product's, bit for bit (RoPE off: the rows are copies)."""
    import torch
    from gpu_harness import Pair, make_qkv
    ref = O.ref_attn_lib()
    assert ref.ref_sizeof_committed_tokens_info() == C.sizeof(B.CommittedTokensInfo)
    rng = np.random.default_rng(23)
    nq, dh, max_requests, max_seq, max_tokens = 8, 64, 6, 256, 128
    H, rows = nq * dh, max_seq + B.MAX_SPEC_TREE_TOKEN_NUM
    p = Pair(B.TREE_VERIFY_MODE, "half", nq, nq, dh, max_requests, max_seq, max_tokens, apply_rope=False, quirks=1)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    kc = torch.zeros(max_requests, rows, H, dtype=torch.float16, device="cuda")
    vc = torch.zeros_like(kc)
    for slot, n in ((1, 100), (4, 33)):
        p.fill_cache(rng, slot, n)
        k, v = p.m.read_kv(slot, 0, n)
        kc[slot, :n], vc[slot, :n] = k.reshape(n, H), v.reshape(n, H)

    def dev(x):
        return torch.from_numpy(np.frombuffer(bytes(x), dtype=np.uint8).copy()).cuda()

    def check(lens):
        for slot, n in lens.items():
            k, v = p.m.read_kv(slot, 0, n)
            torch.cuda.synchronize()
            assert torch.equal(k.reshape(n, H), kc[slot, :n]), slot
            assert torch.equal(v.reshape(n, H), vc[slot, :n]), slot

    def ref_update(bc, qkv_dev):
        ti, ri = dev(bc.tokensInfo), dev(bc.requestsInfo)
        assert ref.ref_update_tree_branch_kv_cache_half(C.c_void_p(qkv_dev.data_ptr()), C.c_void_p(kc.data_ptr()),
                                                        C.c_void_p(vc.data_ptr()), C.c_void_p(ti.data_ptr()),
                                                        C.c_void_p(ri.data_ptr()), dh, bc.num_tokens, rows, H, st) == 0
        torch.cuda.synchronize()

    par_a, par_b = TL.tree_from_widths([1, 1, 3]), TL.random_tree(rng, 12)
    bc1 = TL.make_tree_batch([dict(slot=1, committed_len=100, parent=par_a), dict(slot=4, committed_len=33, parent=par_b)])
    qkv1 = make_qkv(rng, bc1.num_tokens, nq, nq, dh, "half")
    p.gpu_forward(bc1, qkv1)
    q1 = torch.from_numpy(np.ascontiguousarray(qkv1, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    ref_update(bc1, q1)
    check({1: 100 + len(par_a), 4: 33 + len(par_b)})
    # accept root + two nodes of request 1 (batch idx 0, 1, 2) and root + node 1 of request 4 (idx 6, 7); commit them
    off4 = len(par_a)
    commits = [(0, 1, 100), (1, 1, 101), (2, 1, 102), (off4, 4, 33), (off4 + 1, 4, 34)]
    par_c = TL.tree_from_widths([2, 2])
    bc2 = TL.make_tree_batch([dict(slot=1, committed_len=103, parent=par_c), dict(slot=4, committed_len=35, parent=par_c)],
                             commits=commits)
    qkv2 = make_qkv(rng, bc2.num_tokens, nq, nq, dh, "half")
    p.gpu_forward(bc2, qkv2)
    cm = dev(bc2.committed_tokens)
    assert ref.ref_commit_tokens_half(C.c_void_p(q1.data_ptr()), C.c_void_p(kc.data_ptr()), C.c_void_p(vc.data_ptr()),
                                      C.c_void_p(cm.data_ptr()), dh, bc2.num_tokens_to_commit, bc1.num_tokens, rows, H, st) == 0
    q2 = torch.from_numpy(np.ascontiguousarray(qkv2, dtype=np.float32)).to("cuda", torch.float16).contiguous()
    ref_update(bc2, q2)
    check({1: 103 + len(par_c), 4: 35 + len(par_c)})
