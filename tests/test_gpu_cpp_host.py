"""GPU: the C++ host side over the C ABI, compiled and RUN (north_star: "the host side stays in C++").
flexflow_b200/host/incr_decoding.cc - FileDataLoader, RequestManager::prepare_next_batch, (Residual)RMSNorm, Linear,
IncMultiHeadSelfAttention, Softmax, ArgMax and the InferenceResult D2H, all through host/ffb_ops.h - decodes an
attention-only LLaMA-style model from HF-export weight files.  The same model runs through the Python mirror (the
same library calls in the same order), so the token streams must be IDENTICAL; a torch fp32 restatement of the model
pins the first tokens."""
import os
import subprocess
import sys

import numpy as np
import pytest

from flexflow_b200 import batch_config as B

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
L_, HID, NH, NKV, VOCAB, NEW = 3, 256, 4, 2, 512, 24
DH = HID // NH


def _write_weights(folder, rng):
    w = {}

    def put(name, arr):
        w[name] = arr.astype(np.float16)
        w[name].tofile(os.path.join(folder, name))
    put("embed_tokens.weight", rng.standard_normal((VOCAB, HID)))
    for i in range(L_):
        put("layers.%d.input_layernorm.weight" % i, 1 + 0.1 * rng.standard_normal(HID))
        put("layers.%d.self_attn.q_proj.weight" % i, rng.standard_normal((NH * DH, HID)) * 0.08)
        put("layers.%d.self_attn.k_proj.weight" % i, rng.standard_normal((NKV * DH, HID)) * 0.08)
        put("layers.%d.self_attn.v_proj.weight" % i, rng.standard_normal((NKV * DH, HID)) * 0.08)
        put("layers.%d.self_attn.o_proj.weight" % i, rng.standard_normal((HID, NH * DH)) * 0.08)
    put("norm.weight", 1 + 0.1 * rng.standard_normal(HID))
    put("lm_head.weight", rng.standard_normal((VOCAB, HID)) * 0.3)
    return w


def _python_mirror(folder, prompts):
    """The same operator sequence as host/incr_decoding.cc through flexflow_b200.ops / request_manager / file_loader."""
    import torch
    from flexflow_b200 import file_loader as FL
    from flexflow_b200 import ops
    from flexflow_b200.request_manager import RequestManager
    dev, MAXR, MAXT, MAXSEQ, H = "cuda:0", 4, 64, 256, NH * DH
    ld = FL.FileDataLoader(folder, NH, NKV, HID, DH, 1)

    def half(a):
        return torch.from_numpy(np.ascontiguousarray(a).view(np.float16)).to(dev)
    emb = ld.load_from_file("embed_tokens.weight", VOCAB * HID).reshape(VOCAB, HID)
    h = ops.FFHandler(dev)
    rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True)
    layers = []
    for i in range(L_):
        base = "layers.%d" % i
        layers.append(dict(
            nw=half(ld.load_from_file(base + ".input_layernorm.weight", HID)),
            wqkv=half(ld.load_attention_weights_to_dense(base + ".self_attn", False, 0)),
            wo=half(ld.load_attention_weights_to_dense(base + ".self_attn", True, 0)),
            lq=ops.LinearMeta(h, HID, 3 * H, torch.float16, B.DT_NONE, MAXT), lo=ops.LinearMeta(h, H, HID, torch.float16, B.DT_NONE, MAXT),
            at=ops.IncMultiHeadSelfAttentionMeta(h, NH, NKV, DH, torch.float16, rope, False, 1.0, True, False, MAXR, MAXT, MAXSEQ, 0, NH)))
    fnw, head = half(ld.load_from_file("norm.weight", HID)), half(ld.load_from_file("lm_head.weight", VOCAB * HID).reshape(VOCAB, HID))
    lh = ops.LinearMeta(h, HID, VOCAB, torch.float16, B.DT_NONE, MAXT)
    nm = ops.ResidualRMSNormMeta(h, 1e-6, HID)
    buf = lambda *s: torch.empty(*s, dtype=torch.float16, device=dev)
    x, res, xn, qkv, att, o = buf(MAXT, HID), buf(MAXT, HID), buf(MAXT, HID), buf(MAXT, 3 * H), buf(MAXT, H), buf(MAXT, HID)
    logits, probs, idx = buf(MAXT, VOCAB), buf(MAXT, VOCAB), torch.empty(MAXT, dtype=torch.int32, device=dev)
    rm = RequestManager()
    rm.set_max_requests_per_batch(MAXR)
    rm.set_max_tokens_per_batch(MAXT)
    rm.set_max_spec_tree_token_num(0)
    rm.set_max_sequence_length(MAXSEQ)
    guids = [rm.register_new_request(p, max_new_tokens=NEW) for p in prompts]

    def step(bc):
        T = bc.num_tokens
        h.load_batch_config(bc)
        ids = [bc.tokensInfo[t].token_id for t in range(T)]
        x[:T] = half(emb[ids])
        for i, ly in enumerate(layers):
            if i == 0:
                ops.RMSNorm.inference_kernel_wrapper(nm, bc, x[:T], ly["nw"], xn[:T])
            else:
                ops.ResidualRMSNorm.inference_kernel_wrapper(nm, bc, x[:T], o[:T], ly["nw"], res[:T], xn[:T])
                x[:T] = res[:T]
            ops.Linear.inference_kernel_wrapper(ly["lq"], bc, xn, qkv, ly["wqkv"], None, HID, 3 * H, T)
            ops.IncMultiHeadSelfAttention.inference_kernel_wrapper(ly["at"], bc, 0, qkv, att)
            ops.Linear.inference_kernel_wrapper(ly["lo"], bc, att, o, ly["wo"], None, H, HID, T)
        ops.ResidualRMSNorm.inference_kernel_wrapper(nm, bc, x[:T], o[:T], fnw, res[:T], xn[:T])
        ops.Linear.inference_kernel_wrapper(lh, bc, xn, logits, head, None, HID, VOCAB, T)
        ops.Softmax.inference_kernel_wrapper(ops.SoftmaxMeta(h), bc, False, logits[:T], probs[:T])
        ops.ArgMax.forward_kernel_wrapper(ops.ArgMaxMeta(h), bc, probs[:T], idx[:T], None, T)
        ir = B.InferenceResult()
        torch.cuda.synchronize()
        got = idx[:T].cpu().numpy()
        for t in range(T):
            ir.token_ids[t] = int(got[t])
        return ir
    rm.serve_incr_decoding(step)
    return [rm.request_tokens(g) for g in guids]


def _torch_reference(w, prompt, n_new):
    """fp32 restatement of the attention-only model (HF rotate-half RoPE, GQA, causal attention)."""
    import torch
    f = {k: torch.from_numpy(v.astype(np.float32)) for k, v in w.items()}
    toks = list(prompt)
    for _ in range(n_new):
        ids = torch.tensor(toks)
        T = len(toks)
        x = f["embed_tokens.weight"][ids]
        pos = torch.arange(T, dtype=torch.float32)
        inv = 1.0 / (10000.0 ** (torch.arange(0, DH, 2, dtype=torch.float32) / DH))
        ang = pos[:, None] * inv[None, :]
        cos, sin = torch.cat([ang.cos(), ang.cos()], -1), torch.cat([ang.sin(), ang.sin()], -1)
        rot = lambda t: torch.cat([-t[..., DH // 2:], t[..., :DH // 2]], -1)
        for i in range(L_):
            rms = lambda t, wt: t * torch.rsqrt(t.pow(2).mean(-1, keepdim=True) + 1e-6) * wt
            xn = rms(x, f["layers.%d.input_layernorm.weight" % i])
            q = (xn @ f["layers.%d.self_attn.q_proj.weight" % i].t()).view(T, NH, DH)
            k = (xn @ f["layers.%d.self_attn.k_proj.weight" % i].t()).view(T, NKV, DH)
            v = (xn @ f["layers.%d.self_attn.v_proj.weight" % i].t()).view(T, NKV, DH)
            q, k = q * cos[:, None] + rot(q) * sin[:, None], k * cos[:, None] + rot(k) * sin[:, None]
            k, v = k.repeat_interleave(NH // NKV, 1), v.repeat_interleave(NH // NKV, 1)
            s = torch.einsum("qhd,khd->hqk", q, k) / DH ** 0.5
            s = s.masked_fill(torch.triu(torch.ones(T, T, dtype=torch.bool), 1), float("-inf"))
            a = torch.einsum("hqk,khd->qhd", s.softmax(-1), v).reshape(T, NH * DH)
            x = x + a @ f["layers.%d.self_attn.o_proj.weight" % i].t()
        xn = x[-1] * torch.rsqrt(x[-1].pow(2).mean() + 1e-6) * f["norm.weight"]
        toks.append(int((xn @ f["lm_head.weight"].t()).argmax()))
    return toks


def test_cpp_host_program_decodes_the_same_tokens_as_the_python_mirror(tmp_path):
    rng = np.random.default_rng(0)
    folder = str(tmp_path / "weights")
    os.makedirs(folder)
    w = _write_weights(folder, rng)
    prompts = [[int(t) for t in rng.integers(1, VOCAB, size=n)] for n in (7, 19, 3)]
    pfile = str(tmp_path / "prompts.txt")
    with open(pfile, "w") as fh:
        for p in prompts:
            fh.write(" ".join(map(str, p)) + "\n")
    exe = str(tmp_path / "incr_decoding")
    libdir = os.path.join(ROOT, "flexflow_b200")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.run(["g++", "-std=c++17", "-O1", "-I", os.path.join(ROOT, "include"), "-I", os.path.join(libdir, "host"),
                    "-I", os.path.join(cuda, "include"), os.path.join(libdir, "host", "incr_decoding.cc"), "-o", exe,
                    "-L", libdir, "-l:libffb200.so", "-L", os.path.join(cuda, "lib64"), "-lcudart",
                    "-Wl,-rpath," + libdir], check=True)
    r = subprocess.run([exe, folder, str(L_), str(HID), str(NH), str(NKV), str(VOCAB), str(NEW), pfile],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    got = [[int(t) for t in line.split()[2:]] for line in r.stdout.splitlines() if line.startswith("TOKENS")]
    assert len(got) == len(prompts)
    want = _python_mirror(folder, prompts)
    assert got == want
    for p, toks in zip(prompts, got):
        assert toks[:len(p)] == p and len(toks) == len(p) + NEW
        ref = _torch_reference(w, p, 6)
        assert toks[:len(ref)] == ref, (toks[:len(ref)], ref)
