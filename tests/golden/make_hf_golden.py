"""Generates tests/golden/hf_llama_attn.npz from the HuggingFace/PyTorch alignment path the reference's
own tests use (tests/inference/huggingface_inference.py:57-101 for the model call;
tests/inference/inference_alignment_test.py:264-309 for WHICH tensors are compared: FlexFlow's qkv_proj
output vs HF q/k/v_proj outputs, FlexFlow's attention output vs the INPUT of HF o_proj).

Run in the build container (needs transformers + torch CPU):  python tests/golden/make_hf_golden.py
The fixture is committed; the GPU box never runs this script.
"""
import os

import numpy as np
import torch
from transformers import LlamaConfig, LlamaForCausalLM

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hf_llama_attn.npz")


def capture(cfg, seed, prompt_len, steps):
    torch.manual_seed(seed)
    model = LlamaForCausalLM(cfg).eval().float()
    layer = model.model.layers[0].self_attn
    grabbed = {}

    def hook(name):
        def fn(mod, inp, out):
            grabbed.setdefault(name, []).append(out.detach().clone())
        return fn

    def pre_hook(mod, inp):
        grabbed.setdefault("o_in", []).append(inp[0].detach().clone())

    hs = [layer.q_proj.register_forward_hook(hook("q")), layer.k_proj.register_forward_hook(hook("k")),
          layer.v_proj.register_forward_hook(hook("v")), layer.o_proj.register_forward_pre_hook(pre_hook)]
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, cfg.vocab_size, (1, prompt_len), generator=g)
    with torch.no_grad():
        out = model(ids, use_cache=True)
        past = out.past_key_values
        tokens = [int(out.logits[0, -1].argmax())]
        for _ in range(steps - 1):
            out = model(torch.tensor([[tokens[-1]]]), past_key_values=past, use_cache=True)
            past = out.past_key_values
            tokens.append(int(out.logits[0, -1].argmax()))
    for h in hs:
        h.remove()
    nq, nkv, dh = cfg.num_attention_heads, cfg.num_key_value_heads, cfg.hidden_size // cfg.num_attention_heads
    calls = []
    for q, k, v, o in zip(grabbed["q"], grabbed["k"], grabbed["v"], grabbed["o_in"]):
        T = q.shape[1]
        q = q[0].reshape(T, nq, dh)
        # FlexFlow's qkv_proj output carries K/V at q-head granularity (kv heads replicated by the loader,
        # src/runtime/file_loader.cc:293-302)
        k = k[0].reshape(T, nkv, dh).repeat_interleave(nq // nkv, dim=1)
        v = v[0].reshape(T, nkv, dh).repeat_interleave(nq // nkv, dim=1)
        qkv = torch.stack([q.reshape(T, -1), k.reshape(T, -1), v.reshape(T, -1)], dim=1)  # [T][3][H]
        calls.append((qkv.numpy().astype(np.float32), o[0].numpy().astype(np.float32)))
    return calls, ids[0].numpy(), np.array(tokens)


def main():
    blobs = {}
    for name, nq, nkv, dh, theta in (("mha64", 4, 4, 64, 10000.0), ("gqa128", 4, 2, 128, 500000.0)):
        cfg = LlamaConfig(hidden_size=nq * dh, intermediate_size=2 * nq * dh, num_hidden_layers=1,
                          num_attention_heads=nq, num_key_value_heads=nkv, vocab_size=512,
                          max_position_embeddings=256, rope_theta=theta, attention_bias=False)
        calls, ids, toks = capture(cfg, seed=0, prompt_len=12, steps=4)
        blobs[name + "_meta"] = np.array([nq, nkv, dh, 12, len(calls)], dtype=np.int64)
        blobs[name + "_theta"] = np.array([theta], dtype=np.float64)
        for i, (qkv, o) in enumerate(calls):
            blobs["%s_qkv_%d" % (name, i)] = qkv
            blobs["%s_out_%d" % (name, i)] = o
    np.savez_compressed(OUT, **blobs)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
