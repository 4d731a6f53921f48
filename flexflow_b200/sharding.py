"""Tensor-parallel weight layout of the attention block, restated from the reference loader
(src/runtime/file_loader.cc:216-361): shard r of the packed QKV weight holds [Q_r | K_r | V_r] for q heads
r*n_q/TP .. (r+1)*n_q/TP - 1, with each q head's KV head (kv = q_head / (n_q/n_kv)) replicated into the shard at
q-head granularity (:293-302); o_proj is row-parallel — shard r keeps the input columns of its heads (:350-354).
Weights are [out][in] (in fastest), matching Kernels::Linear (linear_kernels.cu:510-528)."""
import torch


def shard_qkv_proj(wq, wk, wv, num_q_heads, num_kv_heads, head_dim, tp, rank):
    """wq [n_q*Dh][hidden], wk/wv [n_kv*Dh][hidden] -> [3*(n_q/tp)*Dh][hidden] for this shard."""
    hq = num_q_heads // tp
    group = num_q_heads // num_kv_heads
    heads = range(rank * hq, (rank + 1) * hq)
    q = torch.cat([wq[h * head_dim:(h + 1) * head_dim] for h in heads], 0)
    k = torch.cat([wk[(h // group) * head_dim:(h // group + 1) * head_dim] for h in heads], 0)
    v = torch.cat([wv[(h // group) * head_dim:(h // group + 1) * head_dim] for h in heads], 0)
    return torch.cat([q, k, v], 0).contiguous()


def shard_o_proj(wo, num_q_heads, head_dim, tp, rank):
    """wo [hidden][n_q*Dh] -> [hidden][(n_q/tp)*Dh]: the columns of this shard's heads."""
    hq = num_q_heads // tp
    return wo[:, rank * hq * head_dim:(rank + 1) * hq * head_dim].contiguous()
