"""CPU: the oracle (C++ restatement of the reference kernels, fp32) in a closed greedy-decoding loop reproduces the
HuggingFace golden tokens (tests/golden/hf_llama160m_greedy.npz, BASELINE.json configs[0]).  Pins the oracle at model
level: RoPE convention, KV placement across steps, prompt + decode attention, projection layout [Q | K | V]."""
import os
import sys

import numpy as np

import ffb_testlib as TL
from flexflow_b200 import batch_config as B
from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
STEPS = 24  # of the 128 golden tokens (CPU time); the GPU test checks all 128


def test_oracle_greedy_matches_hf_golden():
    import torch
    sys.path.insert(0, os.path.join(HERE, "golden"))
    import make_hf_decode_golden as G
    fx = np.load(os.path.join(HERE, "golden", "hf_llama160m_greedy.npz"))
    seed, prompt, want = int(fx["seed"][0]), fx["prompt"], fx["tokens"][:STEPS]
    model = G.build_model(seed)
    cfg = model.config
    nq, dh, hid, eps = cfg.num_attention_heads, cfg.hidden_size // cfg.num_attention_heads, cfg.hidden_size, cfg.rms_norm_eps
    O.set_threads(os.cpu_count() or 1)
    layers = []
    for lyr in model.model.layers:
        a = lyr.self_attn
        layers.append(dict(
            wqkv=torch.cat([a.q_proj.weight, a.k_proj.weight, a.v_proj.weight], 0).detach().numpy(),
            wo=a.o_proj.weight.detach().numpy(),
            attn=O.OracleAttn(B.INC_DECODING_MODE, "float", nq, dh, 2, 64, 0, 32), lyr=lyr))

    def rms(x, w):
        return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w

    def forward(ids, first_depth, is_prompt):
        T = len(ids)
        bc = TL.make_inc_batch([dict(slot=0, first_depth=first_depth, ntok=T, prompt=is_prompt)])
        with torch.no_grad():
            x = model.model.embed_tokens.weight[torch.tensor(ids)]
            for L_ in layers:
                lyr = L_["lyr"]
                xn = rms(x, lyr.input_layernorm.weight).numpy()
                qkv = O.linear(xn, L_["wqkv"], None, "float").reshape(T, 3, hid)
                att = L_["attn"].inc_forward(bc, qkv)
                x = x + torch.from_numpy(O.linear(att, L_["wo"], None, "float"))
                x = x + lyr.mlp(rms(x, lyr.post_attention_layernorm.weight))
            return int((rms(x[-1], model.model.norm.weight) @ model.lm_head.weight.t()).argmax())

    toks = [forward(list(prompt), 0, True)]
    for i in range(STEPS - 1):
        toks.append(forward([toks[-1]], len(prompt) + i, False))
    assert np.array_equal(np.array(toks), want), (toks, want.tolist())
