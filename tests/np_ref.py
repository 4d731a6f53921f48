"""Independent numpy fp64 statement of masked softmax attention + HF-style RoPE, used to cross-check the
C++ oracle (two restatements written separately must agree)."""
import numpy as np


def rope_hf(x, pos, theta=10000.0):
    """x: [heads, dh]; rotate pairs (i, i+dh/2) by pos * theta^(-2i/dh)  (HF rotate_half convention,
    reference apply_rotary_embedding_hf inc_multihead_self_attention.cu:664-738)."""
    dh = x.shape[-1]
    i = np.arange(dh // 2, dtype=np.float64)
    freq = pos / (theta ** (2 * i / dh))
    c, s = np.cos(freq), np.sin(freq)
    a, b = x[..., : dh // 2].astype(np.float64), x[..., dh // 2:].astype(np.float64)
    return np.concatenate([a * c - b * s, a * s + b * c], axis=-1)


def attend(q, K, V, visible, scale):
    """q: [dh]; K, V: [L, dh]; visible: bool[L]."""
    s = (K.astype(np.float64) @ q.astype(np.float64)) * scale
    s = np.where(visible, s, -np.inf)
    if not visible.any():
        return np.zeros(V.shape[1])
    p = np.exp(s - s[visible].max())
    p = np.where(visible, p, 0.0)
    return (p / p.sum()) @ V.astype(np.float64)
