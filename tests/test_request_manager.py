"""CPU: the serving control behind the C ABI (ffb_rm_*, flexflow_b200/csrc/request_manager.cu) against the REFERENCE'S
OWN batch builders compiled into oracle/_ref/libref_host.so (prepare_next_batch, prepare_next_batch_init / _beam /
_verify, store/update_beam_metadata, traverse_beam_tree - src/runtime/request_manager.cc:713-2215,2743-2888), and the
property the reference's own tests pin SpecInfer with: spec_infer output == incr_decoding output
(tests/inference/cpp_inference_tests.sh:185-189).

Both managers are driven in lockstep through serve_spec_infer's schedule (request_manager.cc:3083-3173) over a tiny
transformer whose attention is the oracle's spec / tree-verify / incremental operators (so the KV placement, commit and
mask rules are in the loop); every BeamSearchBatchConfig / TreeVerifyBatchConfig / BatchConfig either side emits is
compared field by field (tests/spec_sim.py:diff_bc), and the final token streams must equal plain greedy decoding."""
import ctypes as C

import numpy as np
import pytest

import spec_sim as S
from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as L

pytestmark = pytest.mark.skipif(S.ref_host_lib() is None, reason="oracle/_ref/libref_host.so not built (make -C oracle ref)")

NEW = 24


def _models(seed, noise=0.15):
    w_llm = S.TinyWeights(seed)
    return w_llm, S.TinyWeights(seed + 100, like=w_llm, noise=noise)


@pytest.mark.parametrize("widths", [[], [1, 1, 3], [3], [2], [1, 3], [1, 2, 1], [3, 1]])
def test_spec_infer_builders_equal_the_reference_and_reproduce_greedy(widths):
    w_llm, w_ssm = _models(1)
    rng = np.random.default_rng(len(widths) + sum(widths))
    prompts = [[int(t) for t in rng.integers(1, w_llm.vocab, size=n)] for n in (5, 9, 3)]
    got, got_ref, stats = S.run_spec_lockstep(w_llm, w_ssm, prompts, [len(p) + NEW for p in prompts], widths)
    assert got == got_ref
    for p, toks in zip(prompts, got):
        want = S.greedy_reference(w_llm, p, NEW)
        assert len(toks) >= len(p) + NEW - 1 and toks == want[:len(toks)], (widths, toks, want)
    # speculation really happened: more than one token per request and round on average, trees wider than a chain
    assert stats["committed"] > stats["rounds"] * len(prompts) * 0.9
    if widths and max(widths) > 1:
        assert stats["max_tree"] > B.MAX_BEAM_DEPTH + 1


@pytest.mark.parametrize("noise", [0.0, 0.6])
def test_spec_infer_with_a_perfect_and_a_poor_ssm(noise):
    """SSM == LLM: every round accepts the whole chain; a poor SSM: most rounds accept only the bonus token."""
    w_llm, w_ssm = _models(7, noise)
    prompts = [[3, 1, 4, 1, 5, 9, 2, 6], [2, 7, 1, 8]]
    got, got_ref, stats = S.run_spec_lockstep(w_llm, w_ssm, prompts, [len(p) + 40 for p in prompts], [1, 1, 3])
    assert got == got_ref
    for p, toks in zip(prompts, got):
        assert toks == S.greedy_reference(w_llm, p, 40)[:len(toks)]
    per_round = stats["committed"] / (stats["rounds"] * len(prompts))
    assert per_round > 6 if noise == 0.0 else per_round < 4, per_round


def test_spec_infer_continuous_batching_more_requests_than_slots():
    """2 slots, 5 requests of different lengths: finished requests free their slot, queued ones load their prompt while
    others speculate (mixed pending / running batches).  Beam width 1 (the reference's ArgTopK takes ONE k per batch)."""
    w_llm, w_ssm = _models(3)
    rng = np.random.default_rng(9)
    prompts = [[int(t) for t in rng.integers(1, w_llm.vocab, size=n)] for n in (4, 11, 6, 2, 8)]
    news = [10, 17, 5, 21, 9]
    got, got_ref, stats = S.run_spec_lockstep(w_llm, w_ssm, prompts, [len(p) + n for p, n in zip(prompts, news)], [],
                                              max_requests=2)
    assert got == got_ref
    for p, n, toks in zip(prompts, news, got):
        assert toks == S.greedy_reference(w_llm, p, n)[:len(toks)] and len(toks) >= len(p) + n - 1


def test_incremental_builder_equals_the_reference_and_reproduces_greedy():
    """serve_incr_decoding's schedule (request_manager.cc:3012-3081): 3 slots, 6 requests, prompts longer than the
    token budget left in a batch (chunked prompt phase), an EOS token that ends some requests early."""
    MAXR, MAXT, SEQ = 3, 16, 128
    w = S.TinyWeights(5)
    rng = np.random.default_rng(2)
    prompts = [[int(t) for t in rng.integers(1, w.vocab, size=n)] for n in (7, 21, 3, 30, 5, 12)]
    news = [12, 6, 20, 4, 9, 15]
    greedy = [S.greedy_reference(w, p, n) for p, n in zip(prompts, news)]
    # pick as EOS a token that some request generates mid-way
    eos = greedy[2][len(prompts[2]) + 7]
    ref = S.RefRequestManager(MAXR, MAXT, 0, SEQ, 0, (), [eos])
    ours = S.make_product_rm(MAXR, MAXT, 0, SEQ, 0, (), [eos])
    guids = []
    for p, n in zip(prompts, news):
        g_ref, g = ref.register_new_request(p, len(p) + n), ours.register_new_request(p, max_length=len(p) + n)
        assert g_ref == g
        guids.append(g)
    llm = S.TinyModel(w, B.INC_DECODING_MODE, MAXR, SEQ, 0, MAXT)
    br, bo, ir = B.BatchConfig().reset(), B.BatchConfig().reset(), B.InferenceResult()
    chunked = False
    for step in range(500):
        br, bo = ref.prepare_next_batch(br, ir), ours.prepare_next_batch(bo, ir)
        assert S.diff_bc("incr", bo, br, MAXR) == [], step
        if bo.num_active_requests() == 0 and ours.num_pending() == 0:
            break
        chunked |= any(bo.requestsInfo[i].prompt_phase and bo.requestsInfo[i].first_token_depth_in_request > 0
                       for i in range(MAXR) if not bo.request_completed[i])
        ir = llm.llm_step(bo)
    else:
        raise AssertionError("incremental decoding did not finish")
    assert chunked, "the scenario was meant to split a prompt over two batches"
    ended_by_eos = 0
    for g, p, n, want in zip(guids, prompts, news, greedy):
        toks = ours.request_tokens(g)
        assert toks == ref.request_tokens(g)
        assert ours.request_status(g) == ref.request_status(g) == 103  # COMPLETED
        gen = want[len(p):]
        if eos in gen[:n - 0]:
            cut = gen.index(eos)
            if len(p) + cut < len(p) + n:
                assert toks == p + gen[:cut], (toks, want)
                ended_by_eos += 1
                continue
        assert toks == want[:len(toks)] and len(toks) == len(p) + n
    assert ended_by_eos >= 1
    assert ours.get_num_processed_requests() == ref.get_num_processed_requests() == len(prompts)
    ours.close()


def test_register_request_length_rules():
    rm = S.make_product_rm(2, 16, 0, 64, 0)
    g = rm.register_new_request([1, 2, 3])                       # both unset: max_sequence_length - 1
    bc = rm.prepare_next_batch(B.BatchConfig().reset(), B.InferenceResult())
    assert bc.requestsInfo[0].request_guid == g and bc.requestsInfo[0].max_length == 63
    g2 = rm.register_new_request([1, 2, 3], max_length=50, max_new_tokens=7)  # max_new_tokens wins
    bc = rm.prepare_next_batch(bc, B.InferenceResult())
    assert bc.requestsInfo[1].request_guid == g2 and bc.requestsInfo[1].max_length == 10
    with pytest.raises(L.FFBError):
        rm.register_new_request(list(range(64)))                 # prompt does not fit
    with pytest.raises(L.FFBError):
        rm.register_new_request([1, 2], max_length=64)           # max_length does not fit
    rm.close()


def test_builders_return_a_status_where_the_reference_asserts():
    rm = S.make_product_rm(2, 16, 8, 64, 1)
    rm.register_new_request([1, 2, 3, 4], max_length=20)
    beam = rm.prepare_next_batch_init(B.TreeVerifyBatchConfig().reset(), B.InferenceResult(), 0)
    assert beam.num_tokens == 4 and not beam.request_running[0]
    bad = B.BeamSearchBatchConfig.from_buffer_copy(beam)
    bad.request_running[0] = True  # a pending request presented as speculating
    with pytest.raises(L.FFBError):
        rm.prepare_next_batch_beam(bad, B.BeamInferenceResult())
    bad = B.BeamSearchBatchConfig.from_buffer_copy(beam)
    bad.requestsInfo[0].request_guid = 77  # unknown request
    with pytest.raises(L.FFBError):
        rm.prepare_next_batch_beam(bad, B.BeamInferenceResult())
    with pytest.raises(L.FFBError):
        rm.prepare_next_batch_init(B.TreeVerifyBatchConfig().reset(), B.InferenceResult(), 3)  # no such SSM
    rm.close()


def test_spec_infer_prompt_longer_than_the_token_budget():
    """A 45-token prompt with max_tokens_per_batch = 16: the SSM loads it over several beam steps (appendPendingRequest),
    the verifier over the verify batch's prompt room, while the second request already speculates."""
    w_llm, w_ssm = _models(1)
    rng = np.random.default_rng(0)
    prompts = [[int(t) for t in rng.integers(1, w_llm.vocab, size=n)] for n in (45, 9)]
    got, got_ref, stats = S.run_spec_lockstep(w_llm, w_ssm, prompts, [len(p) + NEW for p in prompts], [1, 1, 3],
                                              max_requests=2, max_tokens=16, max_spec=16)
    assert got == got_ref
    for p, toks in zip(prompts, got):
        assert toks == S.greedy_reference(w_llm, p, NEW)[:len(toks)] and len(toks) >= len(p) + NEW - 1
