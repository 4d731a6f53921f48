"""CPU: pin the oracle against the HuggingFace alignment path (the reference's own correctness bar:
tests/inference/inference_alignment_test.py:193-204,299-309 — attention output vs o_proj input,
np.allclose(atol=1e-2)).  The fixture was produced by tests/golden/make_hf_golden.py."""
import os

import numpy as np
import pytest

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from oracle import oracle as O

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "hf_llama_attn.npz")


def run_case(name, attn_factory, forward):
    z = np.load(GOLD)
    nq, nkv, dh, prompt_len, ncalls = [int(x) for x in z[name + "_meta"]]
    theta = float(z[name + "_theta"][0])
    attn = attn_factory(nq, nkv, dh, theta)
    depth = 0
    worst = 0.0
    for i in range(ncalls):
        qkv, gold = z["%s_qkv_%d" % (name, i)], z["%s_out_%d" % (name, i)]
        T = qkv.shape[0]
        bc = TL.make_inc_batch([dict(slot=1, first_depth=depth, ntok=T, prompt=(T > 1))])
        out = forward(attn, bc, qkv)
        worst = max(worst, float(np.abs(out - gold).max()))
        depth += T
    return worst


@pytest.mark.parametrize("name", ["mha64", "gqa128"])
def test_oracle_matches_hf_fp32(name):
    def factory(nq, nkv, dh, theta):
        return O.OracleAttn(B.INC_DECODING_MODE, "float", nq, dh, 2, 64, 0, 16, theta=theta)

    worst = run_case(name, factory, lambda a, bc, qkv: a.inc_forward(bc, qkv))
    assert worst < 2e-5, worst  # fp32 vs fp32: only summation order and the 1e-6 softmax epsilon differ
