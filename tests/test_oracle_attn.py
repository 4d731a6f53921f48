"""CPU: the C++ oracle (restatement of the reference kernels) against an independent numpy statement."""
import numpy as np
import pytest

import ffb_testlib as TL
import np_ref
from flexflow_b200 import batch_config as B
from oracle import oracle as O

NQ, DH = 4, 64
H = NQ * DH


def _fill_cache(orc, rng, slot, n, dt):
    orc.kcache[slot, :n] = TL.rand_dt(rng, (n, H), dt)
    orc.vcache[slot, :n] = TL.rand_dt(rng, (n, H), dt)


def _check_token(orc, out, qkv_rot, t, slot, visible, tol=2e-3):
    L = len(visible)
    for h in range(NQ):
        sl = slice(h * DH, (h + 1) * DH)
        ref = np_ref.attend(qkv_rot[t, 0, sl], orc.kcache[slot, :L, sl], orc.vcache[slot, :L, sl],
                            np.asarray(visible), 1.0 / np.sqrt(DH))
        assert np.abs(out[t, sl] - ref).max() < tol, (t, h)


def _rot(qkv, bc):
    r = qkv.astype(np.float64).copy()
    for t in range(bc.num_tokens):
        pos = bc.tokensInfo[t].abs_depth_in_request
        for part in (0, 1):
            r[t, part] = np_ref.rope_hf(qkv[t, part].reshape(NQ, DH), pos).reshape(H)
    return r


@pytest.mark.parametrize("dt", ["float", "half"])
def test_inc_decode_and_prompt(dt):
    rng = np.random.default_rng(0)
    orc = O.OracleAttn(B.INC_DECODING_MODE, dt, NQ, DH, 4, 64, 0, 32)
    _fill_cache(orc, rng, 0, 20, dt)
    _fill_cache(orc, rng, 2, 5, dt)
    bc = TL.make_inc_batch([dict(slot=0, first_depth=20, ntok=1), dict(slot=1, first_depth=0, ntok=6, prompt=True),
                            dict(slot=2, first_depth=5, ntok=3, prompt=True)])
    assert bc.num_tokens == 10 and bc.num_generation_tokens == 1
    qkv = TL.rand_dt(rng, (10, 3, H), dt)
    out = orc.inc_forward(bc, qkv)
    rot = _rot(qkv, bc)
    # KV placement (store_kv_cache): token at abs depth
    assert np.abs(orc.kcache[0, 20] - rot[0, 1]).max() < 2e-3
    assert np.array_equal(orc.vcache[1, 3], qkv[4, 2])
    tol = 2e-3 if dt == "float" else 6e-3
    _check_token(orc, out, rot, 0, 0, [True] * 21, tol)
    for j in range(6):
        _check_token(orc, out, rot, 1 + j, 1, [k <= j for k in range(6)], tol)
    for j in range(3):
        _check_token(orc, out, rot, 7 + j, 2, [k <= 5 + j for k in range(8)], tol)


def test_tree_verify_two_steps_with_commit():
    rng = np.random.default_rng(1)
    dt = "half"
    orc = O.OracleAttn(B.TREE_VERIFY_MODE, dt, NQ, DH, 4, 64, 64, 64)
    _fill_cache(orc, rng, 1, 10, dt)
    parent = TL.tree_from_widths([1, 1, 3])  # 6 nodes
    bc1 = TL.make_tree_batch([dict(slot=1, committed_len=10, parent=parent)])
    qkv1 = TL.rand_dt(rng, (bc1.num_tokens, 3, H), dt)
    out1 = orc.tree_forward(bc1, qkv1)
    rot1 = _rot(qkv1, bc1)
    anc = TL.ancestor_mask(parent)
    for qi in range(len(parent)):
        vis = [True] * 10 + [bool((anc[k] >> qi) & 1) for k in range(len(parent))]
        _check_token(orc, out1, rot1, qi, 1, vis, 6e-3)
    # tree K/V sits at slot = position in batch (update_tree_branch_kv_cache_fused)
    assert np.abs(orc.kcache[1, 10 + 4] - rot1[4, 1]).max() < 2e-3
    # step 2: accept root + nodes 1, 2 (depths 10, 11, 12) -> commit from previous batch indices 0, 1, 2
    parent2 = TL.tree_from_widths([2, 1])
    bc2 = TL.make_tree_batch([dict(slot=1, committed_len=13, parent=parent2)],
                             commits=[(0, 1, 10), (1, 1, 11), (2, 1, 12)])
    qkv2 = TL.rand_dt(rng, (bc2.num_tokens, 3, H), dt)
    orc.tree_forward(bc2, qkv2)
    assert np.abs(orc.kcache[1, 12] - rot1[2, 1]).max() < 2e-3
    assert np.array_equal(orc.vcache[1, 11], qkv1[1, 2])
    # committing an index beyond the previous batch violates the reference assert
    bc3 = TL.make_tree_batch([dict(slot=1, committed_len=13, parent=parent2)], commits=[(40, 1, 10)])
    with pytest.raises(RuntimeError):
        orc.tree_forward(bc3, qkv2)


def test_spec_beam_step():
    rng = np.random.default_rng(2)
    dt = "half"
    orc = O.OracleAttn(B.BEAM_SEARCH_MODE, dt, NQ, DH, 4, 64, 64, 64)
    _fill_cache(orc, rng, 0, 12 + 3, dt)  # 12 committed + root + first layer (1) + second layer (1)
    bc = TL.make_beam_batch([dict(slot=0, non_tree=12, widths=[1, 1, 3])], [dict(slot=1, first_depth=0, ntok=4)])
    bm = bc.causalMask[0]
    assert (bm.non_tree_cache_size, bm.tree_size, bm.prompt_size, bm.this_layer_size) == (12, 6, 1, 3)
    assert bc.num_tokens == 7 and bc.num_generation_tokens == 3 and bc.speculative_request_num == 1
    qkv = TL.rand_dt(rng, (7, 3, H), dt)
    out_lit = orc.spec_forward(bc, qkv)  # literal reference: the prompt request is skipped (quirk bit 1)
    assert np.all(out_lit[3:] == 0)
    orc.set_quirks(1)
    out = orc.spec_forward(bc, qkv)
    assert np.array_equal(out[:3], out_lit[:3])
    rot = _rot(qkv, bc)
    total = 12 + 6 + 1 - 1
    # branch tokens land at prompt+non_tree+tree-1-this_layer + idx = 15, 16, 17
    assert np.abs(orc.kcache[0, 15 + 1] - rot[1, 1]).max() < 2e-3
    for qi in range(3):
        bit = 1 + 6 - 1 - 3 + qi
        vis = [True] * 12 + [bool((bm.mask[k] >> bit) & 1) for k in range(total - 12)]
        # a branch sees root, its ancestors and itself, not its siblings
        assert sum(vis[12:]) == 4
        _check_token(orc, out, rot, qi, 0, vis, 6e-3)
    for j in range(4):
        _check_token(orc, out, rot, 3 + j, 1, [k <= j for k in range(4)], 6e-3)
