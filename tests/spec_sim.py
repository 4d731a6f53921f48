"""Test harness for the serving control (VERDICT G2): the reference's RequestManager functions compiled
into oracle/_ref/libref_host.so behind the same Python face as flexflow_b200.request_manager.RequestManager, a field by
field comparison of batch descriptors, and a tiny transformer (numpy fp32 + the oracle's attention operators) that
plays the LLM and the SSM on the CPU so whole SpecInfer runs can be checked against incremental greedy decoding."""
import ctypes as C
import os

import numpy as np

from flexflow_b200 import batch_config as B
from oracle import oracle as O

_HERE = os.path.dirname(os.path.abspath(__file__))
_REF_HOST_SO = os.path.join(_HERE, "..", "oracle", "_ref", "libref_host.so")
_ref = None


def ref_host_lib():
    """The reference's own host-side serving control (oracle/ref_host_wrap.cpp), or None when it was not built."""
    global _ref
    if _ref is None and os.path.exists(_REF_HOST_SO):
        L = C.CDLL(_REF_HOST_SO)
        vp, i = C.c_void_p, C.c_int
        L.ref_rm_reset.argtypes = [i, i, i, i, i, vp, i, vp, i]
        L.ref_rm_register_request.restype = C.c_size_t
        L.ref_rm_register_request.argtypes = [vp, i, i]
        L.ref_rm_prepare_next_batch.argtypes = [vp, vp, vp]
        L.ref_rm_prepare_next_batch_init.argtypes = [vp, vp, i, vp]
        L.ref_rm_prepare_next_batch_beam.argtypes = [vp, vp, vp]
        L.ref_rm_prepare_next_batch_verify.argtypes = [vp, i, vp]
        L.ref_rm_request_tokens.argtypes = [C.c_size_t, vp, i]
        L.ref_rm_request_status.argtypes = [C.c_size_t]
        L.ref_rm_num_processed.restype = C.c_size_t
        L.ref_rm_decoding_steps.argtypes = [C.c_size_t, vp]
        L.ref_sizeof.restype = C.c_size_t
        L.ref_sizeof.argtypes = [i]
        L.ref_init_bitmask.argtypes = [vp, i]
        L.ref_update_bitmask.argtypes = [vp, i, i]
        L.ref_append_pending.argtypes = [vp, i]
        L.ref_append_bitmask.argtypes = [vp, i, i, i, vp, i, i]
        L.ref_traverse_verify_tree.argtypes = [vp, i, vp, i, vp, i, vp, vp, vp]
        assert L.ref_sizeof(0) == C.sizeof(B.BatchConfig) and L.ref_sizeof(1) == C.sizeof(B.BeamSearchBatchConfig)
        assert L.ref_sizeof(2) == C.sizeof(B.TreeVerifyBatchConfig) and L.ref_sizeof(3) == C.sizeof(B.InferenceResult)
        assert L.ref_sizeof(4) == C.sizeof(B.BeamInferenceResult)
        _ref = L
    return _ref


class RefRequestManager:
    """Same face as flexflow_b200.request_manager.RequestManager over the reference's code (a process-wide singleton:
    creating one resets the previous)."""

    def __init__(self, max_requests, max_tokens, max_spec_tree_tokens, max_seq_len, num_ssms, widths=(), eos=()):
        self.L = ref_host_lib()
        w = (C.c_int * max(1, len(widths)))(*widths)
        e = (C.c_int * max(1, len(eos)))(*eos)
        self.L.ref_rm_reset(max_requests, max_tokens, max_spec_tree_tokens, max_seq_len, num_ssms, w, len(widths), e, len(eos))

    def register_new_request(self, tokens, max_length):
        arr = (C.c_int * len(tokens))(*[int(t) for t in tokens])
        return self.L.ref_rm_register_request(arr, len(tokens), max_length)

    def request_tokens(self, guid):
        arr = (C.c_int * 4096)()
        n = self.L.ref_rm_request_tokens(guid, arr, 4096)
        return list(arr[:n])

    def request_status(self, guid):
        return self.L.ref_rm_request_status(guid)

    def num_pending(self):
        return self.L.ref_rm_pending()

    def get_num_processed_requests(self):
        return self.L.ref_rm_num_processed()

    def prepare_next_batch(self, old_bc, result):
        new_bc = B.BatchConfig()
        self.L.ref_rm_prepare_next_batch(C.addressof(old_bc), C.addressof(result), C.addressof(new_bc))
        return new_bc

    def prepare_next_batch_init(self, old_bc, result, model_id=0):
        new_bc = B.BeamSearchBatchConfig()
        self.L.ref_rm_prepare_next_batch_init(C.addressof(old_bc), C.addressof(result), model_id, C.addressof(new_bc))
        return new_bc

    def prepare_next_batch_beam(self, old_bc, result):
        new_bc = B.BeamSearchBatchConfig()
        self.L.ref_rm_prepare_next_batch_beam(C.addressof(old_bc), C.addressof(result), C.addressof(new_bc))
        return new_bc

    def prepare_next_batch_verify(self, old_batches):
        new_bc = B.TreeVerifyBatchConfig()
        ptrs = (C.c_void_p * len(old_batches))(*[C.addressof(b) for b in old_batches])
        self.L.ref_rm_prepare_next_batch_verify(ptrs, len(old_batches), C.addressof(new_bc))
        return new_bc


# ---- field-by-field comparison of the fields a builder DEFINES (the reference leaves the rest uninitialised) ----
def _req_fields(ri):
    return (ri.first_token_depth_in_request, ri.first_token_offset_in_batch, ri.num_tokens_in_batch, ri.max_length,
            bool(ri.prompt_phase), ri.request_guid)


def diff_bc(kind, a, b, max_requests):
    """kind: 'incr' | 'init' | 'beam' | 'verify'.  Returns a list of human-readable differences (empty = equal)."""
    d = []

    def chk(name, x, y):
        if x != y:
            d.append("%s: %r != %r" % (name, x, y))
    chk("num_tokens", a.num_tokens, b.num_tokens)
    if kind != "verify":
        chk("num_generation_tokens", a.num_generation_tokens, b.num_generation_tokens)
    act_a = [i for i in range(max_requests) if not a.request_completed[i]]
    act_b = [i for i in range(max_requests) if not b.request_completed[i]]
    chk("active slots", act_a, act_b)
    chk("completed beyond max_requests", [bool(a.request_completed[i]) for i in range(max_requests, B.MAX_NUM_REQUESTS)],
        [bool(b.request_completed[i]) for i in range(max_requests, B.MAX_NUM_REQUESTS)])
    for k, i in enumerate(act_a):
        chk("requestsInfo[%d]" % i, _req_fields(a.requestsInfo[i]), _req_fields(b.requestsInfo[i]))
        chk("batch_config_request_id[%d]" % k, a.requestsInfo[k].batch_config_request_id, b.requestsInfo[k].batch_config_request_id)
        if kind != "incr":
            chk("request_running[%d]" % i, bool(a.request_running[i]), bool(b.request_running[i]))
            chk("causalMask[%d]" % i, bytes(a.causalMask[i]), bytes(b.causalMask[i]))
    for t in range(min(a.num_tokens, b.num_tokens)):
        ta, tb = a.tokensInfo[t], b.tokensInfo[t]
        chk("tokensInfo[%d]" % t, (ta.abs_depth_in_request, ta.request_index, ta.token_id),
            (tb.abs_depth_in_request, tb.request_index, tb.token_id))
    if kind in ("init", "beam"):
        chk("model_id", a.model_id, b.model_id)
        if kind == "beam":
            chk("speculative_request_num", a.speculative_request_num, b.speculative_request_num)
        for i in act_a:
            ba, bb = a.beamRequestsInfo[i], b.beamRequestsInfo[i]
            chk("beamRequestsInfo[%d]" % i, (ba.beam_size, ba.current_depth, ba.max_depth, ba.sub_request_num),
                (bb.beam_size, bb.current_depth, bb.max_depth, bb.sub_request_num))
            chk("sub_requests[%d]" % i, a.sub_requests[i], b.sub_requests[i])
            if kind == "beam" and a.request_running[i]:  # the layer's tokens (probs / parents are dead data, see DESIGN)
                n = max(0, min(ba.sub_request_num, 3))
                chk("beam tokens[%d]" % i, list(ba.tokens[:n]), list(bb.tokens[:n]))
                chk("beam probs[%d]" % i, list(ba.probs[:n]), list(bb.probs[:n]))
        for t in range(min(a.num_tokens, b.num_tokens)):
            chk("beamTokenInfo[%d]" % t, a.beamTokenInfo[t].sub_request_index, b.beamTokenInfo[t].sub_request_index)
    if kind == "verify":
        chk("num_tokens_to_commit", a.num_tokens_to_commit, b.num_tokens_to_commit)
        for c in range(min(a.num_tokens_to_commit, b.num_tokens_to_commit)):
            ca, cb = a.committed_tokens[c], b.committed_tokens[c]
            chk("committed_tokens[%d]" % c, (ca.token_index, ca.request_index, ca.token_depth),
                (cb.token_index, cb.request_index, cb.token_depth))
    return d


# ---- a tiny transformer on the CPU: numpy fp32 around the oracle's attention operators ----
class TinyWeights:
    def __init__(self, seed, vocab=61, hidden=32, nq=2, dh=16, layers=2, inter=48, like=None, noise=0.0):
        rng = np.random.default_rng(seed)
        self.vocab, self.hidden, self.nq, self.dh, self.layers = vocab, hidden, nq, dh, layers

        self.emb = rng.standard_normal((vocab, hidden)).astype(np.float32)
        self.wqkv = [rng.standard_normal((3 * nq * dh, hidden)).astype(np.float32) * 0.3 for _ in range(layers)]
        self.wo = [rng.standard_normal((hidden, nq * dh)).astype(np.float32) * 0.3 for _ in range(layers)]
        self.w1 = [rng.standard_normal((inter, hidden)).astype(np.float32) * 0.3 for _ in range(layers)]
        self.w2 = [rng.standard_normal((hidden, inter)).astype(np.float32) * 0.3 for _ in range(layers)]
        self.head = rng.standard_normal((vocab, hidden)).astype(np.float32)
        if like is not None:  # a perturbed copy: agrees with `like` most of the time (the SSM of an LLM)
            for k in ("emb", "head"):
                setattr(self, k, getattr(like, k) + noise * getattr(self, k))
            for k in ("wqkv", "wo", "w1", "w2"):
                setattr(self, k, [a + noise * b for a, b in zip(getattr(like, k), getattr(self, k))])


def _rms(x):
    return x / np.sqrt((x * x).mean(-1, keepdims=True) + 1e-6)


class TinyModel:
    """One model instance = weights + one attention operator state per layer (its KV caches) in one inference mode."""

    def __init__(self, w, mode, max_requests, max_seq, max_spec, max_tokens):
        self.w, self.mode = w, mode
        self.attn = [O.OracleAttn(mode, "float", w.nq, w.dh, max_requests, max_seq, max_spec, max_tokens)
                     for _ in range(w.layers)]
        for a in self.attn:
            a.set_quirks(0)

    def logits(self, bc):
        T = bc.num_tokens
        w = self.w
        if T == 0:
            return np.zeros((0, w.vocab), np.float32)
        ids = np.array([bc.tokensInfo[t].token_id for t in range(T)])
        x = w.emb[ids]
        H = w.nq * w.dh
        for l in range(w.layers):
            qkv = (_rms(x) @ w.wqkv[l].T).reshape(T, 3, H)
            a = self.attn[l]
            if self.mode == B.INC_DECODING_MODE:
                o = a.inc_forward(bc, qkv)
            elif self.mode == B.TREE_VERIFY_MODE:
                o = a.tree_forward(bc, qkv)
            else:
                o = a.spec_forward(bc, qkv)
            x = x + o @ w.wo[l].T
            x = x + np.maximum(_rms(x) @ w.w1[l].T, 0) @ w.w2[l].T
        return (_rms(x) @ w.head.T).astype(np.float32)

    def llm_step(self, bc):
        """ArgMax head -> InferenceResult (inference/models/llama.cc:289-291)."""
        ir = B.InferenceResult()
        lg = self.logits(bc)
        for t in range(bc.num_tokens):
            ir.token_ids[t] = int(np.argmax(lg[t]))
        return ir

    def ssm_step(self, bc, max_requests):
        """softmax -> ArgTopK head in speculative mode -> BeamInferenceResult: ONE k for the whole batch, the
        beam_size of the first active request (src/ops/arg_topk.cu:395-428), candidates of token t at [t*k, t*k+k)."""
        ir = B.BeamInferenceResult()
        lg = self.logits(bc)
        k = 1
        for i in range(max_requests):
            if not bc.request_completed[i]:
                k = bc.beamRequestsInfo[i].beam_size
                break
        for t in range(bc.num_tokens):
            p = np.exp(lg[t] - lg[t].max())
            p /= p.sum()
            order = np.lexsort((np.arange(len(p)), -p))[:k]  # value descending, index ascending
            for j, tok in enumerate(order):
                ir.token_ids[t * k + j] = int(tok)
                ir.probs[t * k + j] = float(p[tok])
        return ir


def greedy_reference(w, prompt, n_new, max_seq=512):
    """Plain incremental greedy decoding of one request with the tiny model (the sequence SpecInfer must reproduce)."""
    m = TinyModel(w, B.INC_DECODING_MODE, 1, max_seq, 0, max(len(prompt), 1) + 1)
    import ffb_testlib as TL
    toks = list(prompt)
    bc = TL.make_inc_batch([dict(slot=0, first_depth=0, ntok=len(prompt), prompt=True)])
    for t, tok in enumerate(prompt):
        bc.tokensInfo[t].token_id = tok
    nxt = int(np.argmax(m.logits(bc)[-1]))
    toks.append(nxt)
    for _ in range(n_new - 1):
        bc = TL.make_inc_batch([dict(slot=0, first_depth=len(toks) - 1, ntok=1)])
        bc.tokensInfo[0].token_id = toks[-1]
        toks.append(int(np.argmax(m.logits(bc)[-1])))
    return toks


def make_product_rm(max_requests, max_tokens, max_spec, max_seq, num_ssms, widths=(), eos=()):
    from flexflow_b200.request_manager import RequestManager
    rm = RequestManager()
    rm.set_max_requests_per_batch(max_requests)
    rm.set_max_tokens_per_batch(max_tokens)
    rm.set_max_spec_tree_token_num(max_spec)
    rm.set_max_sequence_length(max_seq)
    for _ in range(num_ssms):
        rm.register_ssm_model()
    for w in widths:
        rm.push_spec_infer_tree_width(w)
    rm.register_tokenizer(eos)
    return rm


def run_spec_lockstep(w_llm, w_ssm, prompts, max_lengths, widths, max_requests=4, max_tokens=64, max_spec=64, max_seq=256,
                      max_rounds=200):
    """serve_spec_infer (request_manager.cc:3083-3173) driven twice in lockstep - the product's builders and the
    reference's own - over the same tiny LLM / SSM.  Every descriptor either side emits is compared field by field.
    Returns (product token lists, reference token lists, stats)."""
    ref = RefRequestManager(max_requests, max_tokens, max_spec, max_seq, 1, widths)
    ours = make_product_rm(max_requests, max_tokens, max_spec, max_seq, 1, widths)
    guids = []
    for p, ml in zip(prompts, max_lengths):
        g_ref, g = ref.register_new_request(p, ml), ours.register_new_request(p, max_length=ml)
        assert g_ref == g
        guids.append(g)
    cap = max_tokens + max_spec * max_requests
    llm = TinyModel(w_llm, B.TREE_VERIFY_MODE, max_requests, max_seq, max_spec, cap)
    ssm = TinyModel(w_ssm, B.BEAM_SEARCH_MODE, max_requests, max_seq, max_spec, cap)
    tree_r, tree_o, tree_ir = B.TreeVerifyBatchConfig().reset(), B.TreeVerifyBatchConfig().reset(), B.InferenceResult()
    stats = dict(rounds=0, verify_tokens=0, committed=0, max_tree=0)
    for _ in range(max_rounds):
        br, bo = ref.prepare_next_batch_init(tree_r, tree_ir, 0), ours.prepare_next_batch_init(tree_o, tree_ir, 0)
        assert diff_bc("init", bo, br, max_requests) == []
        if br.num_active_requests() == 0 and ref.num_pending() == 0:
            assert ours.num_pending() == 0
            break
        for depth in range(B.MAX_BEAM_DEPTH):
            ir = ssm.ssm_step(bo, max_requests)
            br, bo = ref.prepare_next_batch_beam(br, ir), ours.prepare_next_batch_beam(bo, ir)
            assert diff_bc("beam", bo, br, max_requests) == [], (stats["rounds"], depth)
        tree_r, tree_o = ref.prepare_next_batch_verify([br]), ours.prepare_next_batch_verify([bo])
        assert diff_bc("verify", tree_o, tree_r, max_requests) == [], stats["rounds"]
        tree_ir = llm.llm_step(tree_o)
        stats["rounds"] += 1
        stats["verify_tokens"] += tree_o.num_tokens
        stats["committed"] += tree_o.num_tokens_to_commit
        stats["max_tree"] = max([stats["max_tree"]] + [tree_o.requestsInfo[i].num_tokens_in_batch
                                                      for i in range(max_requests) if tree_o.request_running[i]
                                                      and not tree_o.request_completed[i]])
    else:
        raise AssertionError("SpecInfer did not finish in %d rounds" % max_rounds)
    got_o = [ours.request_tokens(g) for g in guids]
    got_r = [ref.request_tokens(g) for g in guids]
    assert [ours.request_status(g) for g in guids] == [ref.request_status(g) for g in guids]
    assert ours.get_num_processed_requests() == ref.get_num_processed_requests()
    ours.close()
    return got_o, got_r, stats
