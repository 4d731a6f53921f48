"""CPU: the accept / reject walk and the bit-mask builders pinned against the REFERENCE'S OWN CODE.

oracle/_ref/libref_host.so holds RequestManager::{initBitMask, updateBitMask, appendPendingRequest, appendBitMask,
traverse_verify_tree} compiled from /root/reference/src/runtime/request_manager.cc:2382-2517,2583-2741 (oracle/Makefile
target `ref`, text streamed at build time, nothing copied).  Here the product (ffb_*, flexflow_b200/csrc/host_tree.cu),
the oracle's restatement (orc_*, oracle/orc_host.cpp) and that reference code run on the same inputs and must agree bit
for bit: on the reference's tree shapes (spec_infer.cc:337-339 widths (1,1,3); width <= 3 x depth <= 8), on random
trees of up to 31 nodes, and on random accept walks.  From tree index 31 on the reference's 32-bit `1 << j` diverges
from the 64-bit shift the product uses - shown explicitly."""
import ctypes as C

import numpy as np
import pytest

import ffb_testlib as TL
import spec_sim as S
from flexflow_b200 import batch_config as B
from flexflow_b200 import request_tree as RT
from oracle import oracle as O

pytestmark = pytest.mark.skipif(S.ref_host_lib() is None, reason="oracle/_ref/libref_host.so not built (make -C oracle ref)")


def _three():
    return B.BitMask(), B.BitMask(), B.BitMask()


def _append_all(got, orc, ref, nodes, layer_nodes, depth, pre_beam, old_sub):
    arr = (C.c_int * max(1, len(layer_nodes)))(*layer_nodes)
    RT.appendBitMask(got, nodes, layer_nodes, depth)
    assert O.lib().orc_append_bitmask(C.addressof(orc), nodes, arr, depth, 1) == 0
    S.ref_host_lib().ref_append_bitmask(C.addressof(ref), nodes, pre_beam, old_sub, arr, len(layer_nodes), depth)


def _grow(widths, verified=1, non_tree=40):
    """init -> (pending chunk) -> update -> one appendBitMask per beam step, the three implementations side by side.
    Yields after every step."""
    R = S.ref_host_lib()
    got, orc, ref = _three()
    RT.initBitMask(got, 9)
    O.lib().orc_init_bitmask(C.addressof(orc), 9)
    R.ref_init_bitmask(C.addressof(ref), 9)
    yield got, orc, ref
    RT.appendPendingRequest(got, 5)
    O.lib().orc_append_pending_request(C.addressof(orc), 5)
    R.ref_append_pending(C.addressof(ref), 5)
    yield got, orc, ref
    got, orc, ref = _three()  # prepare_next_batch_init starts from a fresh mask (request_manager.cc:1188,1386-1388)
    RT.updateBitMask(got, verified, non_tree)
    O.lib().orc_update_bitmask(C.addressof(orc), verified, non_tree, 1)
    R.ref_update_bitmask(C.addressof(ref), verified, non_tree)
    yield got, orc, ref
    layer_nodes, nodes = [1], 1
    prev_beam = 1
    for d, w in enumerate(widths):
        old_sub = nodes
        nodes *= w
        _append_all(got, orc, ref, nodes, layer_nodes, d + 1, prev_beam, old_sub)
        prev_beam = w
        layer_nodes.append(nodes)
        yield got, orc, ref


REFERENCE_SHAPES = [[1, 1, 3] + [1] * 5, [3] + [1] * 7, [1] * 8, [2, 1, 1], [1, 3], [3], [1, 2, 1, 1, 1, 1, 1, 1]]


@pytest.mark.parametrize("widths", REFERENCE_SHAPES)
def test_bitmask_builders_equal_the_reference_code(widths):
    for step, (got, orc, ref) in enumerate(_grow(widths, verified=3, non_tree=57)):
        assert bytes(got) == bytes(ref), ("product vs reference", widths, step)
        assert bytes(orc) == bytes(ref), ("oracle vs reference", widths, step)
    parent = TL.tree_from_widths(widths)
    assert len(parent) <= 31
    assert [ref.mask[i] for i in range(len(parent))] == TL.ancestor_mask(parent)


def test_bitmask_random_trees_up_to_31_nodes_equal_the_reference_code():
    rng = np.random.default_rng(5)
    done = 0
    while done < 200:
        widths = [int(rng.integers(1, 4)) for _ in range(int(rng.integers(1, 9)))]
        nodes, total, ok = 1, 1, True
        for w in widths:
            nodes *= w
            total += nodes
        if total > 31:
            continue
        done += 1
        for step, (got, orc, ref) in enumerate(_grow(widths, verified=int(rng.integers(1, 9)), non_tree=int(rng.integers(0, 2000)))):
            assert bytes(got) == bytes(ref) and bytes(orc) == bytes(ref), (widths, step)


def test_the_32_bit_shift_of_the_reference_diverges_from_tree_index_31():
    """A 33-node tree (root + 32 leaves under one parent is impossible with width <= 3, so: widths 2,2,2,2 -> 31 nodes,
    then one more layer of 2 x 16 = 32 nodes would overflow 64; use a flat 1 + 40 fan-out instead, which the masks allow
    even though the reference's beam search never builds it).  Nodes 0..30 agree; from node 31 on the reference ORs
    `1 << j` computed in 32-bit int arithmetic (undefined for j >= 31) while the product sets bit j of the 64-bit word."""
    got, orc, ref = _three()
    R = S.ref_host_lib()
    RT.updateBitMask(got, 1, 10)
    R.ref_update_bitmask(C.addressof(ref), 1, 10)
    arr = (C.c_int * 1)(1)
    RT.appendBitMask(got, 40, [1], 1)
    R.ref_append_bitmask(C.addressof(ref), 40, 1, 1, arr, 1, 1)
    for i in range(1, 31):
        assert got.mask[i] == ref.mask[i] == 1 << i
    assert got.mask[0] & ((1 << 31) - 1) == ref.mask[0] & ((1 << 31) - 1)
    want_root = (1 << 41) - 1
    assert got.mask[0] == want_root  # every node sees the root: the intended 64-bit semantics
    assert [got.mask[i] for i in range(31, 41)] == [1 << i for i in range(31, 41)]
    bad = [i for i in range(31, 41) if ref.mask[i] != 1 << i]
    assert bad == list(range(31, 41)) or ref.mask[0] != want_root, "the reference build unexpectedly used 64-bit shifts"


def _ref_traverse(in_tree, out_tree, committed):
    R = S.ref_host_lib()
    flat = lambda xs: (C.c_int * max(1, 2 * len(xs)))(*[v for p in xs for v in p])
    ver = (C.c_int * (2 * len(in_tree) + 2))()
    com = (C.c_int * (2 * len(in_tree) + 2))()
    nc = C.c_int(0)
    nv = R.ref_traverse_verify_tree(flat(in_tree), len(in_tree), flat(out_tree), len(out_tree), flat(committed),
                                    len(committed), ver, com, C.byref(nc))
    return ([(ver[2 * i], ver[2 * i + 1]) for i in range(nv)], [(com[2 * i], com[2 * i + 1]) for i in range(nc.value)])


def test_traverse_verify_tree_reference_cases_equal_the_reference_code():
    in_tree = [(5, 10), (7, 11), (8, 11), (9, 11), (1, 12), (2, 12), (3, 12)]
    out_tree = [(8, 11), (0, 12), (2, 12), (0, 12), (0, 13), (99, 13), (0, 13)]
    committed = [(d, i) for i, (_, d) in enumerate(in_tree)]
    for it, ot in ((in_tree, out_tree), ([(5, 10), (7, 11), (8, 11), (9, 11), (2, 12), (1, 12), (3, 12)], out_tree),
                   (in_tree, [(4, 11)] + out_tree[1:]), (in_tree, out_tree[:4])):
        want = _ref_traverse(it, ot, committed)
        assert RT.traverse_verify_tree(it, ot, committed) == want
        assert O.traverse_verify_tree(it, ot, committed) == want


def test_traverse_verify_tree_random_walks_equal_the_reference_code():
    """Layer-order trees as traverse_beam_tree emits them (width <= 3 per layer, depth <= 8, <= 31 nodes), a small
    vocabulary so that equal (token, depth) pairs also turn up in the wrong slot, truncated LLM outputs (a verify batch
    cut at max_verify_tokens_per_batch), random batch indices."""
    rng = np.random.default_rng(12)
    n_multi = 0
    for _ in range(1500):
        layers = [1] + [int(rng.integers(1, 4)) for _ in range(int(rng.integers(1, 9)))]
        d0 = int(rng.integers(0, 500))
        vocab = int(rng.integers(2, 6))
        in_tree = [(int(rng.integers(0, vocab)), d0 + li) for li, cnt in enumerate(layers) for _ in range(cnt)]
        assert len(in_tree) <= 31
        n_out = len(in_tree) if rng.random() < 0.8 else int(rng.integers(1, len(in_tree) + 1))
        out_tree = [(int(rng.integers(0, vocab)), d + 1) for _, d in in_tree[:n_out]]
        committed = [(d, int(rng.integers(0, 1024))) for _, d in in_tree]
        want = _ref_traverse(in_tree, out_tree, committed)
        assert RT.traverse_verify_tree(in_tree, out_tree, committed) == want
        assert O.traverse_verify_tree(in_tree, out_tree, committed) == want
        n_multi += len(want[0]) > 2
    assert n_multi > 100  # the walks really accept beyond the root
