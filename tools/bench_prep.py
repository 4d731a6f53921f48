"""prep + attention operator at configs[3] shapes (8 x 64 tree tokens over 1024 committed), CUDA-graph replay of 8 layers'
operators back to back: us per operator.  Used to compare prep_kernel launch shapes (FFB_PREP_THREADS, ...)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ffb_testlib as TL  # noqa: E402
from flexflow_b200 import batch_config as B  # noqa: E402
from flexflow_b200 import ops  # noqa: E402


def main():
    nq, dh, nreq, ntree, committed, layers = 32, 128, 8, 64, 1024, 8
    h = ops.FFHandler("cuda:0")
    rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True)
    metas = [ops.TreeIncMultiHeadSelfAttentionMeta(h, nq, nq, dh, rotary_embedding_meta=rope, max_requests_per_batch=nreq,
                                                   max_tokens_per_batch=nreq * ntree, max_sequence_length=committed + ntree + 16)
             for _ in range(layers)]
    rng = np.random.default_rng(0)
    g = torch.Generator(device="cuda").manual_seed(0)
    for m in metas:
        for s in range(nreq):
            kv = torch.randn(committed, nq, dh, generator=g, device="cuda", dtype=torch.float16)
            m.write_kv(s, 0, kv, kv)
    bc = TL.make_tree_batch([dict(slot=s, committed_len=committed, parent=TL.random_tree(rng, ntree)) for s in range(nreq)])
    T = bc.num_tokens
    qkv = torch.randn(T, 3, nq * dh, generator=g, device="cuda", dtype=torch.float16)
    out = torch.empty(T, nq * dh, device="cuda", dtype=torch.float16)
    h.load_batch_config(bc)
    bc.num_tokens_to_commit = 0

    def step():
        for m in metas:
            ops.TreeIncMultiHeadSelfAttention.inference_kernel_wrapper(m, bc, 0, qkv, out)
    step()
    torch.cuda.synchronize()
    st = torch.cuda.Stream()
    with torch.cuda.stream(st):
        step()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            step()
        for _ in range(3):
            graph.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            graph.replay()
        e1.record()
    torch.cuda.synchronize()
    print({"op_us": e0.elapsed_time(e1) * 1e3 / 20 / layers, "env": {k: v for k, v in os.environ.items() if k.startswith("FFB_")}}, flush=True)


if __name__ == "__main__":
    main()
