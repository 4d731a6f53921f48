"""Synthetic batch builders restating what RequestManager::prepare_next_batch* emit
(src/runtime/request_manager.cc:713-1138, 1610-1892, 1923-2215), shared by CPU and GPU tests."""
import ctypes as C

import numpy as np

from flexflow_b200 import batch_config as B
from oracle import oracle as O


def _activate(bc, slot, k):
    bc.request_completed[slot] = False
    bc.request_running[slot] = True
    bc.requestsInfo[k].batch_config_request_id = slot  # compaction table (request_manager.cc:859)
    bc.requestsInfo[slot].request_guid = 1000000 + slot
    bc.requestsInfo[slot].max_length = 1 << 20


def make_inc_batch(reqs):
    """reqs: list of dict(slot, first_depth, ntok, prompt). Generation requests must come first
    (the reference's generation kernel indexes tokens by active-request order, inc_…cu:413-435)."""
    bc = B.BatchConfig().reset()
    ngen = 0
    for k, r in enumerate(sorted(reqs, key=lambda r: r["slot"])):
        s = r["slot"]
        _activate(bc, s, k)
        ri = bc.requestsInfo[s]
        ri.first_token_depth_in_request = r["first_depth"]
        ri.first_token_offset_in_batch = bc.num_tokens
        ri.num_tokens_in_batch = r["ntok"]
        ri.prompt_phase = bool(r.get("prompt", False))
        if not ri.prompt_phase:
            assert r["ntok"] == 1
            ngen += 1
        for j in range(r["ntok"]):
            ti = bc.tokensInfo[bc.num_tokens]
            ti.request_index, ti.abs_depth_in_request, ti.token_id = s, r["first_depth"] + j, 7 + j
            bc.num_tokens += 1
    bc.num_generation_tokens = ngen
    return bc


def tree_from_widths(widths):
    """Layer-order tree like the reference builds (beam widths per depth, spec_infer.cc:337-339):
    returns parent[] with node 0 the root."""
    parent, layer = [-1], [0]
    for w in widths:
        nxt = []
        for p in layer:
            for _ in range(w):
                parent.append(p)
                nxt.append(len(parent) - 1)
        layer = nxt
    return parent


def random_tree(rng, n):
    """Random tree in layer (BFS) order: node i's parent is an earlier node, depths non-decreasing."""
    parent, depth = [-1], [0]
    while len(parent) < n:
        # attach to a node of the deepest or second-deepest layer to keep BFS order valid
        cands = [i for i, d in enumerate(depth) if d >= max(depth) - 1]
        p = int(rng.choice(cands))
        parent.append(p)
        depth.append(depth[p] + 1)
    order = sorted(range(n), key=lambda i: (depth[i], i))
    remap = {o: i for i, o in enumerate(order)}
    return [(-1 if parent[o] < 0 else remap[parent[o]]) for o in order]


def tree_depths(parent):
    d = []
    for i, p in enumerate(parent):
        d.append(0 if p < 0 else d[p] + 1)
    return d


def ancestor_mask(parent):
    """mask[k] bit q  <=>  k is q or an ancestor of q (what appendBitMask builds, request_manager.cc:2448-2517)."""
    n = len(parent)
    mask = [0] * n
    for q in range(n):
        k = q
        while k >= 0:
            mask[k] |= 1 << q
            k = parent[k]
    return mask


def make_tree_batch(reqs, commits=()):
    """reqs: list of dict(slot, committed_len, parent=[...]) for running requests (tree verify) or
    dict(slot, first_depth, ntok, prompt=True) for prompt chunks.  commits: (token_index, slot, depth)."""
    bc = B.TreeVerifyBatchConfig().reset()
    for k, r in enumerate(sorted(reqs, key=lambda r: r["slot"])):
        s = r["slot"]
        _activate(bc, s, k)
        ri = bc.requestsInfo[s]
        ri.first_token_offset_in_batch = bc.num_tokens
        if r.get("prompt", False):
            ri.first_token_depth_in_request = r["first_depth"]
            ri.num_tokens_in_batch = r["ntok"]
            ri.prompt_phase = True
            for j in range(r["ntok"]):
                ti = bc.tokensInfo[bc.num_tokens]
                ti.request_index, ti.abs_depth_in_request, ti.token_id = s, r["first_depth"] + j, 11 + j
                bc.num_tokens += 1
            bm = bc.causalMask[s]
            bm.non_tree_cache_size, bm.tree_size, bm.prompt_size, bm.this_layer_size = 0, 1, r["first_depth"] + r["ntok"], r["ntok"]
        else:
            parent = r["parent"]
            depth = tree_depths(parent)
            ri.first_token_depth_in_request = r["committed_len"]
            ri.num_tokens_in_batch = len(parent)
            ri.prompt_phase = False
            bm = bc.causalMask[s]
            bm.non_tree_cache_size = r["committed_len"]
            bm.tree_size, bm.prompt_size, bm.this_layer_size = len(parent), 1, 1
            for i, m in enumerate(ancestor_mask(parent)):
                bm.mask[i] = m
            for j in range(len(parent)):
                ti = bc.tokensInfo[bc.num_tokens]
                ti.request_index, ti.abs_depth_in_request, ti.token_id = s, r["committed_len"] + depth[j], 13 + j
                bc.num_tokens += 1
    for (tok, slot, dep) in commits:
        c = bc.committed_tokens[bc.num_tokens_to_commit]
        c.token_index, c.request_index, c.token_depth = tok, slot, dep
        bc.num_tokens_to_commit += 1
    return bc


def make_beam_batch(spec_reqs, prompt_reqs=()):
    """spec_reqs: dict(slot, non_tree, widths=[w1, w2, ...]) — the SSM has already expanded
    len(widths)-1 layers and is now processing the last one (sub_request_num = nodes of that layer).
    prompt_reqs: dict(slot, first_depth, ntok).  Speculative requests are emitted first
    (request_manager.cc:1639-1763)."""
    bc = B.BeamSearchBatchConfig().reset()
    k = 0
    for r in sorted(spec_reqs, key=lambda r: r["slot"]):
        s = r["slot"]
        _activate(bc, s, k)
        k += 1
        bm = bc.causalMask[s]
        O.lib().orc_update_bitmask(C.addressof(bm), 1, r["non_tree"], 1)
        layer_nodes = [1]
        nodes = 1
        for d, w in enumerate(r["widths"]):
            nodes *= w
            arr = (C.c_int * len(layer_nodes))(*layer_nodes)
            rc = O.lib().orc_append_bitmask(C.addressof(bm), nodes, arr, d + 1, 1)
            assert rc == 0, rc
            layer_nodes.append(nodes)
        ri = bc.requestsInfo[s]
        # first_token_depth = processed tokens; all branch tokens share abs_depth (request_manager.cc:1743-1757)
        ri.first_token_depth_in_request = r["non_tree"] + len(r["widths"])
        ri.first_token_offset_in_batch = bc.num_tokens
        ri.num_tokens_in_batch = 1
        ri.prompt_phase = False  # new_bc is default-constructed for running requests (request_manager.cc:1655-1763)
        bri = bc.beamRequestsInfo[s]
        bri.sub_request_num, bri.beam_size, bri.current_depth = nodes, 1, len(r["widths"])
        bc.sub_requests[s] = nodes
        for b in range(nodes):
            ti = bc.tokensInfo[bc.num_tokens]
            ti.request_index, ti.abs_depth_in_request, ti.token_id = s, ri.first_token_depth_in_request, 17 + b
            bc.beamTokenInfo[bc.num_tokens].sub_request_index = b
            bc.num_tokens += 1
            bc.num_generation_tokens += 1
    bc.speculative_request_num = k
    for r in sorted(prompt_reqs, key=lambda r: r["slot"]):
        s = r["slot"]
        _activate(bc, s, k)
        k += 1
        bm = bc.causalMask[s]
        O.lib().orc_init_bitmask(C.addressof(bm), r["first_depth"] + r["ntok"])
        bm.this_layer_size = r["ntok"]
        ri = bc.requestsInfo[s]
        ri.first_token_depth_in_request = r["first_depth"]
        ri.first_token_offset_in_batch = bc.num_tokens
        ri.num_tokens_in_batch = r["ntok"]
        ri.prompt_phase = True
        bc.sub_requests[s] = 1
        bc.beamRequestsInfo[s].sub_request_num = 1
        for j in range(r["ntok"]):
            ti = bc.tokensInfo[bc.num_tokens]
            ti.request_index, ti.abs_depth_in_request, ti.token_id = s, r["first_depth"] + j, 19 + j
            bc.num_tokens += 1
    return bc


def rand_dt(rng, shape, dt, scale=1.0):
    return O.round_to_dt((rng.standard_normal(shape) * scale).astype(np.float32), dt)


def expand_kv(x, nq, nkv):
    """[..., nkv, dh] -> [..., nq, dh] replicating kv heads (file_loader.cc:293-302 layout)."""
    return np.repeat(x, nq // nkv, axis=-2)


def rel_err(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-12))
