"""GPU parity: CUDA attention operators (through the C ABI) vs the CPU oracle on identical inputs.
Tolerance: 1e-2 relative (north_star) for fp16/bf16 outputs; KV placement bit-exact for V and within
one DT ulp of the oracle's float RoPE for K."""
import os

import numpy as np
import pytest

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as ffb_lib  # noqa: E402

pytestmark = pytest.mark.gpu
TOL = 1e-2


def _pair(*a, **k):
    from gpu_harness import Pair
    return Pair(*a, **k)


def _qkv(*a, **k):
    from gpu_harness import make_qkv
    return make_qkv(*a, **k)


@pytest.mark.parametrize("dt,nq,nkv,dh", [("half", 8, 8, 128), ("half", 12, 12, 64), ("bf16", 8, 8, 128),
                                          ("half", 8, 2, 128), ("bf16", 8, 1, 64)])
@pytest.mark.parametrize("force", [None, "decode", "multi"])
def test_inc_decode(dt, nq, nkv, dh, force, monkeypatch):
    if force:
        monkeypatch.setenv("FFB_FORCE_ATTN", force)
        ffb_lib.reload_env()
    rng = np.random.default_rng(10)
    p = _pair(B.INC_DECODING_MODE, dt, nq, nkv, dh, 8, 700, 64)
    lens = {0: 1, 2: 77, 3: 300, 5: 689, 6: 64}
    for s, n in lens.items():
        p.fill_cache(rng, s, n)
    bc = TL.make_inc_batch([dict(slot=s, first_depth=n, ntok=1) for s, n in lens.items()])
    qkv = _qkv(rng, bc.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc, qkv)
    assert np.isfinite(got).all()
    assert TL.rel_err(got, want) < TOL
    # appended K/V rows
    for s, n in lens.items():
        gk, gv = p.gpu_cache(s, n + 1)
        ok, ov = p.oracle_cache(s, n + 1)
        assert np.array_equal(gv, ov)
        assert np.abs(gk - ok).max() <= 2e-2 * max(1.0, np.abs(ok).max())
        assert TL.rel_err(gk, ok) < 5e-3


@pytest.mark.parametrize("dt,nq,nkv,dh", [("half", 4, 4, 128), ("half", 4, 4, 64), ("bf16", 4, 2, 128)])
def test_inc_prompt_and_mixed(dt, nq, nkv, dh):
    rng = np.random.default_rng(11)
    p = _pair(B.INC_DECODING_MODE, dt, nq, nkv, dh, 6, 512, 512)
    p.fill_cache(rng, 0, 200)
    p.fill_cache(rng, 3, 37)
    # decode request, fresh prompt (new == total), chunked prompt continuing at depth 37 (> 64 rows: 2 m-tiles)
    bc = TL.make_inc_batch([dict(slot=0, first_depth=200, ntok=1), dict(slot=2, first_depth=0, ntok=70, prompt=True),
                            dict(slot=3, first_depth=37, ntok=130, prompt=True)])
    qkv = _qkv(rng, bc.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc, qkv)
    assert np.isfinite(got).all()
    assert TL.rel_err(got, want) < TOL


def test_inc_long_context_split_kv():
    """Contexts beyond the reference's 48 KiB shared-memory logits limit."""
    rng = np.random.default_rng(12)
    p = _pair(B.INC_DECODING_MODE, "half", 4, 4, 128, 2, 9000, 16, with_oracle=True)
    p.fill_cache(rng, 0, 8191)
    p.fill_cache(rng, 1, 4000)
    bc = TL.make_inc_batch([dict(slot=0, first_depth=8191, ntok=1), dict(slot=1, first_depth=4000, ntok=1)])
    qkv = _qkv(rng, 2, 4, 4, 128, "half")
    got, want = p.forward_both(bc, qkv)
    assert TL.rel_err(got, want) < TOL


@pytest.mark.parametrize("dt,nq,nkv,dh", [("half", 8, 8, 128), ("half", 12, 12, 64), ("bf16", 8, 2, 128)])
def test_tree_verify_steps(dt, nq, nkv, dh):
    rng = np.random.default_rng(13)
    p = _pair(B.TREE_VERIFY_MODE, dt, nq, nkv, dh, 6, 600, 256)
    p.fill_cache(rng, 1, 100)
    p.fill_cache(rng, 4, 513)
    par_a = TL.tree_from_widths([1, 1, 3])          # reference-shaped tree (spec_infer.cc:337-339), 6 nodes
    par_b = TL.tree_from_widths([3, 1, 1, 1, 1, 1, 1, 1])  # 1 + 3*8 = 25 nodes: the reference's maximum
    bc1 = TL.make_tree_batch([dict(slot=1, committed_len=100, parent=par_a),
                              dict(slot=2, first_depth=0, ntok=20, prompt=True),
                              dict(slot=4, committed_len=513, parent=par_b)])
    qkv1 = _qkv(rng, bc1.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc1, qkv1)
    assert np.isfinite(got).all()
    assert TL.rel_err(got, want) < TOL
    # step 2: commit root+2 accepted nodes of request 1 (batch idx 0,1,2), the prompt of request 2
    # (idx 6..25 -> depths 0..19) and root+1 of request 4 (idx 26, 27)
    commits = [(0, 1, 100), (1, 1, 101), (2, 1, 102)] + [(6 + j, 2, j) for j in range(20)] + [(26, 4, 513), (27, 4, 514)]
    par_c = TL.tree_from_widths([2, 2, 1])
    bc2 = TL.make_tree_batch([dict(slot=1, committed_len=103, parent=par_c),
                              dict(slot=2, committed_len=20, parent=par_a),
                              dict(slot=4, committed_len=515, parent=par_c)], commits=commits)
    qkv2 = _qkv(rng, bc2.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc2, qkv2)
    assert TL.rel_err(got, want) < TOL
    for s, n in ((1, 103), (2, 20), (4, 515)):
        gk, gv = p.gpu_cache(s, n)
        ok, ov = p.oracle_cache(s, n)
        assert np.array_equal(gv, ov)
        assert TL.rel_err(gk, ok) < 5e-3


def test_tree_verify_64_nodes_wide_mask():
    """64 tree tokens per request (BASELINE config 4) needs the 1ULL mask semantics."""
    rng = np.random.default_rng(14)
    p = _pair(B.TREE_VERIFY_MODE, "half", 8, 8, 128, 8, 1100, 512)
    reqs = []
    for s in range(8):
        p.fill_cache(rng, s, 1024 - s * 3)
        reqs.append(dict(slot=s, committed_len=1024 - s * 3, parent=TL.random_tree(rng, 64)))
    bc = TL.make_tree_batch(reqs)
    assert bc.num_tokens == 512
    qkv = _qkv(rng, 512, 8, 8, 128, "half")
    got, want = p.forward_both(bc, qkv)
    assert TL.rel_err(got, want) < TOL


@pytest.mark.parametrize("tc", ["1", "0"])
@pytest.mark.parametrize("dt,nq,nkv,dh", [("half", 32, 8, 128), ("bf16", 16, 16, 64)])
def test_tree_verify_ragged_stream_k(dt, nq, nkv, dh, tc, monkeypatch):
    """Ragged request lengths x many heads: the tcgen05 kernel's stream-K schedule cuts units into 3+ pieces
    (split records + last-arriver merge), several row tiles per request (GQA), a prompt chunk in the same batch.
    tc=0 runs the same batch through the mma.sync kernel."""
    monkeypatch.setenv("FFB_ATTN_TC", tc)
    ffb_lib.reload_env()
    rng = np.random.default_rng(21)
    p = _pair(B.TREE_VERIFY_MODE, dt, nq, nkv, dh, 8, 3100, 512)
    lens = {0: 3000, 1: 50, 3: 700, 4: 1500, 6: 129, 7: 2047}
    reqs = []
    for i, (s, n) in enumerate(lens.items()):
        p.fill_cache(rng, s, n)
        reqs.append(dict(slot=s, committed_len=n, parent=TL.random_tree(rng, 16 + 9 * i)))
    reqs.insert(2, dict(slot=2, first_depth=0, ntok=150, prompt=True))
    bc = TL.make_tree_batch(reqs)
    qkv = _qkv(rng, bc.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc, qkv)
    assert np.isfinite(got).all()
    assert TL.rel_err(got, want) < TOL
    # per-token check: no row may be off (a lost or doubly counted split piece would hit single rows)
    per_tok = np.abs(got - want).reshape(bc.num_tokens, -1).max(axis=1) / max(1e-6, np.abs(want).max())
    assert per_tok.max() < 2e-2, (int(per_tok.argmax()), float(per_tok.max()))


def test_tree_commit_state_error():
    from flexflow_b200.lib import FFBError
    rng = np.random.default_rng(15)
    p = _pair(B.TREE_VERIFY_MODE, "half", 4, 4, 64, 4, 128, 64, with_oracle=False)
    bc = TL.make_tree_batch([dict(slot=0, committed_len=4, parent=[-1, 0])], commits=[(5, 0, 3)])
    with pytest.raises(FFBError):
        p.gpu_forward(bc, _qkv(rng, 2, 4, 4, 64, "half"))


@pytest.mark.parametrize("dt,nq,nkv,dh", [("half", 12, 12, 64), ("half", 8, 8, 128)])
def test_spec_beam_steps(dt, nq, nkv, dh):
    rng = np.random.default_rng(16)
    p = _pair(B.BEAM_SEARCH_MODE, dt, nq, nkv, dh, 6, 400, 128, quirks=1)
    p.fill_cache(rng, 0, 50 + 3)
    p.fill_cache(rng, 3, 300 + 2)
    bc = TL.make_beam_batch([dict(slot=0, non_tree=50, widths=[1, 1, 3]), dict(slot=3, non_tree=300, widths=[3, 1])],
                            [dict(slot=4, first_depth=0, ntok=9)])
    qkv = _qkv(rng, bc.num_tokens, nq, nkv, dh, dt)
    got, want = p.forward_both(bc, qkv)
    assert np.isfinite(got).all()
    assert TL.rel_err(got, want) < TOL


@pytest.mark.parametrize("name", ["mha64", "gqa128"])
def test_hf_golden_fp16(name):
    """The committed HuggingFace fixture through the CUDA path (fp16): reference bar atol=1e-2."""
    from test_oracle_golden import run_case

    def factory(nq, nkv, dh, theta):
        return _pair(B.INC_DECODING_MODE, "half", nq, nkv, dh, 2, 64, 16, theta=theta, with_oracle=False)

    worst = run_case(name, factory, lambda a, bc, qkv: a.gpu_forward(bc, qkv))
    assert worst < 1e-2, worst
