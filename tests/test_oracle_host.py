"""CPU: integer host logic (bit masks, tree accept/commit) and the quantisation format in the oracle."""
import ctypes as C

import numpy as np

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from oracle import oracle as O


def _append(bm, new_nodes, layer_nodes, depth, wide=1):
    arr = (C.c_int * len(layer_nodes))(*layer_nodes)
    return O.lib().orc_append_bitmask(C.addressof(bm), new_nodes, arr, depth, wide)


def test_append_bitmask_equals_ancestor_closure():
    # widths (1,1,3) as spec_infer.cc:337-339, then a wider (3,1,1) tree
    for widths in ([1, 1, 3], [3, 1, 1], [2, 2, 1], [3] + [1] * 7):
        bm = B.BitMask()
        O.lib().orc_update_bitmask(C.addressof(bm), 1, 40, 1)
        assert (bm.non_tree_cache_size, bm.tree_size, bm.prompt_size) == (40, 1, 1)
        layer_nodes, nodes = [1], 1
        for d, w in enumerate(widths):
            nodes *= w
            assert _append(bm, nodes, layer_nodes, d + 1) == 0
            layer_nodes.append(nodes)
        parent = TL.tree_from_widths(widths)
        assert bm.tree_size == len(parent)
        assert [bm.mask[i] for i in range(len(parent))] == TL.ancestor_mask(parent)


def test_bitmask_32bit_quirk_matches_wide_for_small_trees():
    """the reference shifts a 32-bit 1; identical to 1ULL for tree index <= 30."""
    a, b = B.BitMask(), B.BitMask()
    for bm, wide in ((a, 0), (b, 1)):
        O.lib().orc_update_bitmask(C.addressof(bm), 1, 7, wide)
        layer_nodes, nodes = [1], 1
        for d, w in enumerate([3] + [1] * 7):  # 25 nodes, the reference's maximum
            nodes *= w
            assert _append(bm, nodes, layer_nodes, d + 1, wide) == 0
            layer_nodes.append(nodes)
    assert bytes(a) == bytes(b)


def test_append_bitmask_rejects_more_than_64_nodes():
    bm = B.BitMask()
    O.lib().orc_update_bitmask(C.addressof(bm), 1, 0, 1)
    assert _append(bm, 64, [1], 1) == 1


def test_traverse_verify_tree_accepts_one_path():
    # input tree (layer order): root(5,d10); layer: (7,d11) (8,d11) (9,d11); layer: (1,d12) (2,d12) (3,d12)
    in_tree = [(5, 10), (7, 11), (8, 11), (9, 11), (1, 12), (2, 12), (3, 12)]
    # LLM outputs per node (token, depth+1): root predicts 8 -> node 2 accepted; node 2 (slot 1) predicts 2 ->
    # node (2,d12) is in slot 1 of its layer -> accepted; it predicts 99
    out_tree = [(8, 11), (0, 12), (2, 12), (0, 12), (0, 13), (99, 13), (0, 13)]
    committed = [(d, i) for i, (_, d) in enumerate(in_tree)]
    ver, new_c = O.traverse_verify_tree(in_tree, out_tree, committed)
    assert ver == [(8, 11), (2, 12), (99, 13)]
    assert new_c == [(10, 0), (11, 2), (12, 5)]
    # wrong layer slot: same token in a different branch is NOT accepted
    in2 = [(5, 10), (7, 11), (8, 11), (9, 11), (2, 12), (1, 12), (3, 12)]
    ver2, new_c2 = O.traverse_verify_tree(in2, out_tree, committed)
    assert ver2 == [(8, 11), (2, 12)] and new_c2 == [(10, 0), (11, 2)]
    # nothing accepted beyond the root
    ver3, new_c3 = O.traverse_verify_tree(in_tree, [(4, 11)] + out_tree[1:], committed)
    assert ver3 == [(4, 11)] and new_c3 == [(10, 0)]


def test_quant_blob_format():
    rng = np.random.default_rng(3)
    in_dim, out_dim = 64, 64
    w = (rng.standard_normal((out_dim, in_dim)) * 0.02).astype(np.float32)
    for bits in (4, 8):
        blob = O.quantize(w, bits)
        assert blob.size == O.blob_bytes(bits, in_dim, out_dim)
        deq = O.dequant(blob, bits, in_dim, out_dim).view(np.float16).astype(np.float32)
        assert np.abs(deq - w).max() < 0.2 / (2 ** bits - 1)
    # int4 nibble order: even element index = high nibble (decompress_kernels.cu:43-50)
    blob = O.quantize(w, 4)
    nvals = in_dim * out_dim // 2
    offs = blob[nvals:nvals + 2 * (in_dim * out_dim // 32)].view(np.float16).astype(np.float32).reshape(out_dim // 32, in_dim)
    scales = blob[nvals + 2 * (in_dim * out_dim // 32):].view(np.float16).astype(np.float32).reshape(out_dim // 32, in_dim)
    o, i = 5, 10  # even flat index
    q_hi = blob[(o * in_dim + i) // 2] >> 4
    deq = O.dequant(blob, 4, in_dim, out_dim).view(np.float16).astype(np.float32)
    assert abs(deq[o, i] - (q_hi / scales[0, i] + offs[0, i])) < 1e-3
