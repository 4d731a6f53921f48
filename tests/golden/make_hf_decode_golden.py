"""Generates tests/golden/hf_llama160m_greedy.npz: BASELINE.json configs[0] — LLaMA-160M architecture
(JackFram/llama-160m shapes, random init from a fixed seed: there is no network for checkpoints), fp32, one prompt of
16 token ids, 128 greedy tokens — through the HuggingFace/PyTorch path the reference's alignment tests use
(tests/inference/huggingface_inference.py:57-101: `generate(do_sample=False)` == argmax of the last position's logits
with a KV cache).  Stores the prompt, the generated ids and, per step, the gap between the two largest logits in
units of the logit standard deviation (how fragile that step's argmax is to 16-bit arithmetic).

Run in the build container (transformers + torch CPU):  python tests/golden/make_hf_decode_golden.py [seed]
The fixture is committed; the GPU test rebuilds the same weights from the seed and never reads /root/reference.
"""
import os
import sys

import numpy as np
import torch
from transformers import LlamaConfig, LlamaForCausalLM

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hf_llama160m_greedy.npz")
PROMPT_LEN, NEW_TOKENS = 16, 128


def llama160m_config():
    return LlamaConfig(hidden_size=768, intermediate_size=3072, num_hidden_layers=12, num_attention_heads=12,
                       num_key_value_heads=12, vocab_size=32000, max_position_embeddings=2048, rope_theta=10000.0,
                       rms_norm_eps=1e-6, attention_bias=False, tie_word_embeddings=False)


def build_model(seed):
    torch.manual_seed(seed)
    return LlamaForCausalLM(llama160m_config()).eval().float()


def prompt_ids(seed):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(3, 32000, (1, PROMPT_LEN), generator=g)


def greedy(model, ids, steps):
    tokens, margins = [], []
    with torch.no_grad():
        out = model(ids, use_cache=True)
        for _ in range(steps):
            logits = out.logits[0, -1].double()
            top = torch.topk(logits, 2).values
            margins.append(float((top[0] - top[1]) / logits.std()))
            tokens.append(int(logits.argmax()))
            out = model(torch.tensor([[tokens[-1]]]), past_key_values=out.past_key_values, use_cache=True)
    return np.array(tokens, dtype=np.int64), np.array(margins)


def main():
    seed = int(sys.argv[1]) if len(sys.argv) > 1 else 0
    model = build_model(seed)
    ids = prompt_ids(seed)
    tokens, margins = greedy(model, ids, NEW_TOKENS)
    print("seed", seed, "min margin %.4f sigma at step %d; tokens[:8] %s" % (margins.min(), int(margins.argmin()), tokens[:8]))
    if "--dry" not in sys.argv:
        np.savez_compressed(OUT, seed=np.array([seed]), prompt=ids[0].numpy(), tokens=tokens, margins=margins,
                            torch_version=np.array([torch.__version__]))
        print("wrote", OUT)


if __name__ == "__main__":
    main()
