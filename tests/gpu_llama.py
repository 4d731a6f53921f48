"""GPU test helper: a LLaMA-architecture model whose attention hot path (qkv_proj -> {Inc | SpecInc | TreeInc}
MultiHeadSelfAttention -> o_proj) and token selection (ArgMax / ArgTopK + the InferenceResult D2H) run through the C ABI,
in any of the three inference modes, driven by a batch descriptor exactly as InferenceManager::inference drives the
reference's operators (src/runtime/inference_manager.cc:408-468).  The operators outside the path (embedding, the
norms, the MLP, lm_head: inference/models/llama.cc:60-93,198-276) are plain torch fp32 on the same GPU with an fp32
residual stream, as in tests/test_gpu_greedy_decode.py."""
import ctypes as C

import torch

from flexflow_b200 import batch_config as B
from flexflow_b200 import lib as L
from flexflow_b200 import ops


def _rms(x, w, eps):
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * w


class PathModel:
    ATTN = {B.INC_DECODING_MODE: (ops.IncMultiHeadSelfAttentionMeta, ops.IncMultiHeadSelfAttention),
            B.BEAM_SEARCH_MODE: (ops.SpecIncMultiHeadSelfAttentionMeta, ops.SpecIncMultiHeadSelfAttention),
            B.TREE_VERIFY_MODE: (ops.TreeIncMultiHeadSelfAttentionMeta, ops.TreeIncMultiHeadSelfAttention)}

    def __init__(self, hf_model, mode, dtype=torch.float16, max_requests=4, max_tokens=64, max_seq=512, max_spec=64,
                 dev="cuda:0", weight_noise=0.0, noise_seed=0):
        cfg = hf_model.config
        self.mode, self.dtype, self.dev = mode, dtype, dev
        self.nq = cfg.num_attention_heads
        self.hid = cfg.hidden_size
        self.dh = self.hid // self.nq
        self.eps = cfg.rms_norm_eps
        self.vocab = cfg.vocab_size
        self.handler = ops.FFHandler(dev)
        self.cap = max_tokens + (max_spec * max_requests if mode != B.INC_DECODING_MODE else 0)
        rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=float(getattr(cfg, "rope_theta", 10000.0)))
        meta_cls, self.attn_op = self.ATTN[mode]
        g = torch.Generator(device="cpu").manual_seed(noise_seed)

        def w(t):  # the SSM of the closed-loop test is a perturbed copy of the LLM
            t = t.detach().float()
            if weight_noise:
                t = t + weight_noise * t.std() * torch.randn(t.shape, generator=g)
            return t
        self.layers = []
        for lyr in hf_model.model.layers:
            a = lyr.self_attn
            wqkv = torch.cat([w(a.q_proj.weight), w(a.k_proj.weight), w(a.v_proj.weight)], 0)
            self.layers.append(dict(
                wqkv=wqkv.to(dev, dtype).contiguous(), wo=w(a.o_proj.weight).to(dev, dtype).contiguous(),
                lin_qkv=ops.LinearMeta(self.handler, self.hid, 3 * self.hid, dtype=dtype, max_tokens_per_batch=self.cap),
                lin_o=ops.LinearMeta(self.handler, self.hid, self.hid, dtype=dtype, max_tokens_per_batch=self.cap),
                attn=meta_cls(self.handler, self.nq, self.nq, self.dh, dtype, rope, False, 1.0, True, False, max_requests,
                              self.cap, max_seq, max_spec, self.nq),
                ln1=w(lyr.input_layernorm.weight).to(dev), ln2=w(lyr.post_attention_layernorm.weight).to(dev),
                gate=w(lyr.mlp.gate_proj.weight).to(dev), up=w(lyr.mlp.up_proj.weight).to(dev),
                down=w(lyr.mlp.down_proj.weight).to(dev)))
        self.emb = w(hf_model.model.embed_tokens.weight).to(dev)
        self.norm_w, self.head = w(hf_model.model.norm.weight).to(dev), w(hf_model.lm_head.weight).to(dev)
        self.qkv_buf = torch.empty(self.cap, 3 * self.hid, dtype=dtype, device=dev)
        self.att_buf = torch.empty(self.cap, self.hid, dtype=dtype, device=dev)
        self.o_buf = torch.empty(self.cap, self.hid, dtype=dtype, device=dev)
        self.idx_buf = torch.empty(self.cap * B.MAX_SPECULATIVE_TREE_BRANCHES, dtype=torch.int32, device=dev)
        self.prob_buf = torch.empty(self.cap * B.MAX_SPECULATIVE_TREE_BRANCHES, dtype=torch.float32, device=dev)
        self.last_logits = None

    def _stream(self):
        return C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def logits(self, bc):
        """One model step over the batch `bc`: fp32 logits [num_tokens][vocab]."""
        T = bc.num_tokens
        assert 0 < T <= self.cap
        self.handler.load_batch_config(bc)
        ids = torch.tensor([bc.tokensInfo[t].token_id for t in range(T)], device=self.dev)
        x = self.emb[ids]
        hid = self.hid
        for L_ in self.layers:
            xn = _rms(x, L_["ln1"], self.eps).to(self.dtype).contiguous()
            ops.Linear.inference_kernel_wrapper(L_["lin_qkv"], bc, xn, self.qkv_buf, L_["wqkv"], None, hid, 3 * hid, T)
            self.attn_op.inference_kernel_wrapper(L_["attn"], bc, 0, self.qkv_buf, self.att_buf)
            ops.Linear.inference_kernel_wrapper(L_["lin_o"], bc, self.att_buf, self.o_buf, L_["wo"], None, hid, hid, T)
            x = x + self.o_buf[:T].float()
            xn = _rms(x, L_["ln2"], self.eps)
            x = x + (torch.nn.functional.silu(xn @ L_["gate"].t()) * (xn @ L_["up"].t())) @ L_["down"].t()
        self.last_logits = _rms(x, self.norm_w, self.eps) @ self.head.t()
        return self.last_logits

    def llm_step(self, bc):
        """ArgMax head + InferenceResult D2H (src/ops/argmax.cc:300-372) through ffb_argmax / ffb_fetch_inference_result."""
        ir = B.InferenceResult()
        T = bc.num_tokens
        if T == 0:
            return ir
        lg = self.logits(bc).to(self.dtype).contiguous()
        ops.ArgMax.forward_kernel_wrapper(ops.ArgMaxMeta(self.handler), bc, lg, self.idx_buf[:T], None, T)
        L.check(L.load().ffb_fetch_inference_result(self.idx_buf.data_ptr(), T, C.addressof(ir), self._stream()),
                "ffb_fetch_inference_result")
        return ir

    def ssm_step(self, bc, max_requests):
        """softmax -> ArgTopK head in speculative mode + BeamInferenceResult D2H (inference/models/llama.cc:277-284,
        src/ops/arg_topk.cc:407-437): ONE k per batch, the first active request's beam_size (arg_topk.cu:395-428)."""
        ir = B.BeamInferenceResult()
        T = bc.num_tokens
        if T == 0:
            return ir
        k = next(bc.beamRequestsInfo[i].beam_size for i in range(max_requests) if not bc.request_completed[i])
        lg = self.logits(bc).to(self.dtype).contiguous()
        probs = torch.empty_like(lg)
        ops.Softmax.inference_kernel_wrapper(ops.SoftmaxMeta(self.handler), bc, False, lg, probs)
        ops.ArgTopK.forward_kernel_wrapper(ops.ArgTopKMeta(self.handler, k, True), bc, probs, self.prob_buf[:T * k],
                                           self.idx_buf[:T * k], T)
        L.check(L.load().ffb_fetch_beam_inference_result(self.idx_buf.data_ptr(), self.prob_buf.data_ptr(), T, k,
                                                         C.addressof(ir), self._stream()),
                "ffb_fetch_beam_inference_result")
        return ir

    def close(self):
        for L_ in self.layers:
            L_["attn"].close()
            L_["lin_qkv"].close()
            L_["lin_o"].close()
