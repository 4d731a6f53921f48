"""CPU: the product's host-side tree logic (C ABI, flexflow_b200/csrc/host_tree.cu) is bit-exact against the oracle's
restatement of RequestManager::{init,update,append}BitMask / traverse_verify_tree on the reference's tree shapes and
on random trees (no CUDA call is made: the library only has to load)."""
import ctypes as C

import numpy as np
import pytest

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as L
from flexflow_b200 import request_tree as RT
from oracle import oracle as O


def _orc_append(bm, new_nodes, layer_nodes, depth):
    arr = (C.c_int * max(1, len(layer_nodes)))(*layer_nodes)
    return O.lib().orc_append_bitmask(C.addressof(bm), new_nodes, arr, depth, 1)


def _same(a, b):
    return bytes(a) == bytes(b)


@pytest.mark.parametrize("widths", [[1, 1, 3], [3, 1, 1], [2, 2, 1], [3] + [1] * 7, [2, 2, 2, 2, 2], [63], [1] * 20])
def test_bitmask_sequence_matches_oracle(widths):
    got, want = B.BitMask(), B.BitMask()
    RT.initBitMask(got, 9)
    O.lib().orc_init_bitmask(C.addressof(want), 9)
    assert _same(got, want)
    RT.appendPendingRequest(got, 5)
    O.lib().orc_append_pending_request(C.addressof(want), 5)
    assert _same(got, want)
    RT.updateBitMask(got, 1, 40)
    O.lib().orc_update_bitmask(C.addressof(want), 1, 40, 1)
    assert _same(got, want)
    layer_nodes, nodes = [1], 1
    for d, w in enumerate(widths):
        nodes *= w
        RT.appendBitMask(got, nodes, layer_nodes, d + 1)
        assert _orc_append(want, nodes, layer_nodes, d + 1) == 0
        assert _same(got, want), (widths, d)
        layer_nodes.append(nodes)
    parent = TL.tree_from_widths(widths)
    assert [got.mask[i] for i in range(len(parent))] == TL.ancestor_mask(parent)


def test_bitmask_append_random_layers_match_oracle():
    """Arbitrary (also uneven) layer widths, multi-token prompts, dirty masks: same words, same verdict."""
    rng = np.random.default_rng(11)
    for _ in range(300):
        got, want = B.BitMask(), B.BitMask()
        for i in range(B.MAX_SPEC_TREE_TOKEN_NUM):
            got.mask[i] = want.mask[i] = int(rng.integers(0, 2 ** 63)) if rng.random() < 0.3 else 0
        prompt, ntcs = int(rng.integers(1, 4)), int(rng.integers(0, 100))
        for bm in (got, want):
            bm.prompt_size, bm.tree_size, bm.non_tree_cache_size = prompt, prompt, ntcs
        layer_nodes = [prompt]
        for d in range(1, int(rng.integers(2, 8))):
            new = int(rng.integers(1, 12))
            rc_want = _orc_append(want, new, layer_nodes, d)
            rc_got = L.load().ffb_bitmask_append(C.addressof(got), new, (C.c_int * len(layer_nodes))(*layer_nodes), d)
            assert (rc_got == 0) == (rc_want == 0)
            if rc_want:
                break
            assert _same(got, want)
            layer_nodes.append(new)


def test_bitmask_append_rejects_more_than_64_nodes():
    bm = B.BitMask()
    RT.updateBitMask(bm, 1, 0)
    with pytest.raises(L.FFBError):
        RT.appendBitMask(bm, 64, [1], 1)


def test_traverse_verify_tree_reference_cases():
    in_tree = [(5, 10), (7, 11), (8, 11), (9, 11), (1, 12), (2, 12), (3, 12)]
    out_tree = [(8, 11), (0, 12), (2, 12), (0, 12), (0, 13), (99, 13), (0, 13)]
    committed = [(d, i) for i, (_, d) in enumerate(in_tree)]
    assert RT.traverse_verify_tree(in_tree, out_tree, committed) == ([(8, 11), (2, 12), (99, 13)], [(10, 0), (11, 2), (12, 5)])
    in2 = [(5, 10), (7, 11), (8, 11), (9, 11), (2, 12), (1, 12), (3, 12)]
    assert RT.traverse_verify_tree(in2, out_tree, committed) == ([(8, 11), (2, 12)], [(10, 0), (11, 2)])
    assert RT.traverse_verify_tree(in_tree, [(4, 11)] + out_tree[1:], committed) == ([(4, 11)], [(10, 0)])
    with pytest.raises(L.FFBError):  # committed depth disagrees with the tree (reference assert)
        RT.traverse_verify_tree(in_tree, out_tree, [(d + 1, i) for d, i in committed])
    with pytest.raises(L.FFBError):  # more outputs than tree nodes
        RT.traverse_verify_tree(in_tree[:3], out_tree, committed)


def test_traverse_verify_tree_random_trees_match_oracle():
    rng = np.random.default_rng(12)
    for _ in range(500):
        widths = [int(rng.integers(1, 4)) for _ in range(int(rng.integers(1, 6)))]
        layers, n = [1], 1
        for w in widths:
            n = min(n * w, 16)
            layers.append(n)
        d0 = int(rng.integers(0, 50))
        vocab = int(rng.integers(2, 6))  # small vocabulary: many (token, depth) matches, also in the wrong slot
        in_tree = [(int(rng.integers(0, vocab)), d0 + li) for li, cnt in enumerate(layers) for _ in range(cnt)][:64]
        n_out = len(in_tree) if rng.random() < 0.8 else int(rng.integers(1, len(in_tree) + 1))
        out_tree = [(int(rng.integers(0, vocab)), d + 1) for _, d in in_tree[:n_out]]
        committed = [(d, int(rng.integers(0, 512))) for _, d in in_tree]
        assert RT.traverse_verify_tree(in_tree, out_tree, committed) == O.traverse_verify_tree(in_tree, out_tree, committed)
