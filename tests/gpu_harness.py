"""GPU side of the parity tests: drives libffb200.so through flexflow_b200.ops (C ABI) with torch tensors
as device buffers and mirrors every step into the CPU oracle."""
import numpy as np
import torch

from flexflow_b200 import batch_config as B
from flexflow_b200 import ops
from oracle import oracle as O

TORCH_DT = {"half": torch.float16, "bf16": torch.bfloat16}
META = {B.INC_DECODING_MODE: (ops.IncMultiHeadSelfAttentionMeta, ops.IncMultiHeadSelfAttention, "inc_forward"),
        B.TREE_VERIFY_MODE: (ops.TreeIncMultiHeadSelfAttentionMeta, ops.TreeIncMultiHeadSelfAttention, "tree_forward"),
        B.BEAM_SEARCH_MODE: (ops.SpecIncMultiHeadSelfAttentionMeta, ops.SpecIncMultiHeadSelfAttention, "spec_forward")}


class Pair:
    """One attention operator instantiated twice: CUDA (product) and oracle (checker)."""

    def __init__(self, mode, dt, nq, nkv, dh, max_requests, max_seq, max_tokens, theta=10000.0, apply_rope=True,
                 llama3=False, with_oracle=True, quirks=1):
        self.mode, self.dt, self.nq, self.nkv, self.dh = mode, dt, nq, nkv, dh
        self.H = nq * dh
        self.handler = ops.FFHandler("cuda:0")
        meta_cls, self.op, self.orc_fn = META[mode]
        rm = ops.RotaryEmbeddingMeta(apply_rotary_embedding=apply_rope, rope_theta=theta,
                                     rope_type="llama3" if llama3 else "default")
        self.m = meta_cls(self.handler, nq, nkv, dh, dtype=TORCH_DT[dt], rotary_embedding_meta=rm,
                          max_requests_per_batch=max_requests, max_tokens_per_batch=max_tokens,
                          max_sequence_length=max_seq)
        self.orc = None
        if with_oracle:
            self.orc = O.OracleAttn(mode, dt, nq, dh, max_requests, max_seq, B.MAX_SPEC_TREE_TOKEN_NUM, max_tokens,
                                    apply_rope=apply_rope, theta=theta, llama3=llama3)
            self.orc.set_quirks(quirks)

    def fill_cache(self, rng, slot, n, scale=1.0):
        """Random committed K/V rows [0, n) for a request, identical in both implementations."""
        from ffb_testlib import rand_dt, expand_kv
        k = rand_dt(rng, (n, self.nkv, self.dh), self.dt, scale)
        v = rand_dt(rng, (n, self.nkv, self.dh), self.dt, scale)
        self.m.write_kv(slot, 0, torch.from_numpy(k).to("cuda", TORCH_DT[self.dt]),
                        torch.from_numpy(v).to("cuda", TORCH_DT[self.dt]))
        if self.orc is not None:
            self.orc.kcache[slot, :n] = expand_kv(k, self.nq, self.nkv).reshape(n, self.H)
            self.orc.vcache[slot, :n] = expand_kv(v, self.nq, self.nkv).reshape(n, self.H)

    def gpu_forward(self, bc, qkv):
        T = bc.num_tokens
        x = torch.from_numpy(np.ascontiguousarray(qkv, dtype=np.float32)).to("cuda", TORCH_DT[self.dt]).contiguous()
        out = torch.full((max(T, 1), self.H), float("nan"), dtype=TORCH_DT[self.dt], device="cuda")
        self.handler.load_batch_config(bc)
        self.op.inference_kernel_wrapper(self.m, bc, 0, x, out)
        torch.cuda.synchronize()
        return out[:T].float().cpu().numpy()

    def forward_both(self, bc, qkv):
        got = self.gpu_forward(bc, qkv)
        want = getattr(self.orc, self.orc_fn)(bc, qkv)
        return got, want

    def gpu_cache(self, slot, n):
        k, v = self.m.read_kv(slot, 0, n)
        torch.cuda.synchronize()
        return k.float().cpu().numpy(), v.float().cpu().numpy()

    def oracle_cache(self, slot, n):
        G = self.nq // self.nkv
        k = self.orc.kcache[slot, :n].reshape(n, self.nq, self.dh)[:, ::G]
        v = self.orc.vcache[slot, :n].reshape(n, self.nq, self.dh)[:, ::G]
        return k, v


def make_qkv(rng, T, nq, nkv, dh, dt, scale=1.0):
    """[T][3][H] with K/V at q-head granularity, kv heads replicated (file_loader.cc:293-302)."""
    from ffb_testlib import rand_dt, expand_kv
    q = rand_dt(rng, (T, nq, dh), dt, scale)
    k = expand_kv(rand_dt(rng, (T, nkv, dh), dt, scale), nq, nkv)
    v = expand_kv(rand_dt(rng, (T, nkv, dh), dt, scale), nq, nkv)
    return np.stack([q.reshape(T, -1), k.reshape(T, -1), v.reshape(T, -1)], axis=1)
