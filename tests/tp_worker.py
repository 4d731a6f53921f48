"""Worker for test_gpu_tp.py / the gloo CPU test: builds one attention block with FFModel at TP = world size and
compares with the unsharded oracle-free reference computed by rank 0 at TP = 1 on the same weights."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import ffb_testlib as TL  # noqa: E402
from flexflow_b200 import ops  # noqa: E402
from flexflow_b200.model import FFConfig, FFModel  # noqa: E402
from flexflow_b200.sharding import shard_o_proj, shard_qkv_proj  # noqa: E402
from flexflow_b200 import lib as ffb_lib  # noqa: E402

HIDDEN, NQ, NKV, DH = (int(x) for x in os.environ.get("TP_DIMS", "1024,8,8,128").split(","))


def build(tp, rank, comm, dev):
    cfg = FFConfig(tensor_parallelism_degree=tp, tp_rank=rank, max_requests_per_batch=4, max_tokens_per_batch=64,
                   max_sequence_length=256, device=dev)
    m = FFModel(cfg, comm)
    x = m.create_tensor(HIDDEN)
    qkv = m.dense(x, 3 * (NQ // tp) * DH, name="layers.0.self_attn.qkv_proj")
    a = m.inc_multiquery_self_attention(qkv, HIDDEN, NQ, NKV, rotary_embedding_meta=ops.RotaryEmbeddingMeta(True),
                                        name="layers.0.self_attn")
    m.dense(a, HIDDEN, name="layers.0.self_attn.o_proj")
    m.compile_inference()
    return m, x


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(rank)
    dev = "cuda:%d" % rank
    dist.init_process_group("nccl", device_id=torch.device(dev))

    def bcast(b):
        o = [b]
        dist.broadcast_object_list(o, src=0)
        return o[0]
    comm = ops.Communicator(rank, world, bcast)
    g = torch.Generator().manual_seed(5)
    wq = (torch.randn(NQ * DH, HIDDEN, generator=g) * 0.03).half()
    wk = (torch.randn(NKV * DH, HIDDEN, generator=g) * 0.03).half()
    wv = (torch.randn(NKV * DH, HIDDEN, generator=g) * 0.03).half()
    wo = (torch.randn(HIDDEN, NQ * DH, generator=g) * 0.03).half()
    hidden = torch.randn(12, HIDDEN, generator=g).half()
    bc = TL.make_inc_batch([dict(slot=0, first_depth=0, ntok=9, prompt=True), dict(slot=2, first_depth=0, ntok=3, prompt=True)])
    outs = []
    for tp, r, c in ((world, rank, comm), (1, 0, None)):
        m, x = build(tp, r, c, dev)
        m.set_weight("layers.0.self_attn.qkv_proj", shard_qkv_proj(wq, wk, wv, NQ, NKV, DH, tp, r).to(dev))
        m.set_weight("layers.0.self_attn.o_proj", shard_o_proj(wo, NQ, DH, tp, r).to(dev))
        buf = torch.zeros(64, HIDDEN, dtype=torch.float16, device=dev)
        buf[:12] = hidden.to(dev)
        y = m.inference(bc, x, buf)
        # decode one more token per request on top of the prompt KV
        bc2 = TL.make_inc_batch([dict(slot=0, first_depth=9, ntok=1), dict(slot=2, first_depth=3, ntok=1)])
        y2 = m.inference(bc2, x, buf)
        torch.cuda.synchronize()
        outs.append((y.float().cpu().numpy().copy(), y2.float().cpu().numpy().copy()))
    for a, b in zip(outs[0], outs[1]):
        err = np.abs(a - b).max() / max(np.abs(b).max(), 1e-6)
        assert err < 1e-2, err
    if os.environ.get("TP_EXPECT_PEER"):
        # the decode step above went through the fused o_proj + NVLink AllReduce kernel: check the peer mapping exists,
        # that 40 further collectives (epoch parity, flag reuse) keep matching the linear + NCCL path, and that every
        # rank holds bit-identical results (fixed-order fp32 reduction)
        assert comm.has_peer, "communicator has no NVLink peer mapping"
        m, x = build(world, rank, comm, dev)
        m.set_weight("layers.0.self_attn.qkv_proj", shard_qkv_proj(wq, wk, wv, NQ, NKV, DH, world, rank).to(dev))
        m.set_weight("layers.0.self_attn.o_proj", shard_o_proj(wo, NQ, DH, world, rank).to(dev))
        buf = torch.zeros(64, HIDDEN, dtype=torch.float16, device=dev)
        worst = 0.0
        for step in range(40):
            ntok = 1 + step % 3
            buf[:16] = torch.randn(16, HIDDEN, generator=g).half().to(dev)
            bcs = TL.make_inc_batch([dict(slot=s, first_depth=step, ntok=1) for s in range(1 + step % 4)])
            del ntok
            os.environ.pop("FFB_DISABLE_FUSED_ALLREDUCE", None)
            ffb_lib.reload_env()
            y_f = m.inference(bcs, x, buf).float().clone()
            os.environ["FFB_DISABLE_FUSED_ALLREDUCE"] = "1"
            ffb_lib.reload_env()
            for lyr in m.layers:  # same KV state for the second pass: it only re-appends the same rows
                pass
            y_n = m.inference(bcs, x, buf).float().clone()
            torch.cuda.synchronize()
            worst = max(worst, float((y_f - y_n).abs().max() / y_n.abs().max().clamp_min(1e-6)))
            gathered = [torch.empty_like(y_f) for _ in range(world)]
            dist.all_gather(gathered, y_f)
            assert all(torch.equal(gathered[0], t) for t in gathered), "ranks disagree on the fused result"
        os.environ.pop("FFB_DISABLE_FUSED_ALLREDUCE", None)
        ffb_lib.reload_env()
        assert worst < 5e-3, worst
        # large messages: the two-shot peer-memory allreduce kernel (peer_allreduce.cu) behind ffb_allreduce.  The
        # result must be the fp32 sum of the ranks' 16-bit inputs in RANK ORDER rounded once - bit for bit, on every
        # rank, in place and out of place, over many back-to-back collectives (epoch parity, flag reuse) and sizes
        # (whole slices, ragged last slice / chunk, the 16 MiB maximum).
        import ctypes as C
        from flexflow_b200 import batch_config as B
        from flexflow_b200 import lib as L
        sizes = [8, 4096, 12 * 1024, 512 * 4096, 1000 * 1000, 8 * 1024 * 1024, 64 * 4096 + 8, 4 * 1024 * 1024]
        for it in range(30):
            n = sizes[it % len(sizes)]
            dt = torch.float16 if it % 2 == 0 else torch.bfloat16
            xin = (torch.randn(n, generator=g) * 2).to(dt).to(dev) * (rank + 1)
            parts = [torch.empty_like(xin) for _ in range(world)]
            dist.all_gather(parts, xin)
            want = parts[0].float()
            for t in parts[1:]:
                want = want + t.float()
            want = want.to(dt)
            y = xin.clone() if it % 3 == 0 else torch.empty_like(xin)
            src = y if it % 3 == 0 else xin
            st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
            L.check(L.load().ffb_allreduce(comm._h, C.c_void_p(src.data_ptr()), C.c_void_p(y.data_ptr()), n,
                                           B.DT_HALF if dt == torch.float16 else B.DT_BF16, st), "ffb_allreduce")
            torch.cuda.synchronize()
            assert torch.equal(y.view(torch.int16), want.view(torch.int16)), "peer allreduce differs (n=%d, it=%d)" % (n, it)
        # back-to-back collectives with no host synchronisation in between (slot reuse across epochs): x <- sum over
        # ranks, 6 times in place; every rank must hold the same bits as a torch replay of the recurrence
        n = 512 * 4096
        xin = (torch.randn(n, generator=g) * 0.5).half().to(dev) * (rank + 1)
        parts = [torch.empty_like(xin) for _ in range(world)]
        dist.all_gather(parts, xin)
        y = xin.clone()
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
        for _ in range(6):
            L.check(L.load().ffb_allreduce(comm._h, C.c_void_p(y.data_ptr()), C.c_void_p(y.data_ptr()), n, B.DT_HALF, st), "ffb_allreduce")
        torch.cuda.synchronize()
        want = parts[0].float()
        for t in parts[1:]:
            want = want + t.float()
        want = want.half()
        for _ in range(5):
            acc = want.float()
            for _r in range(1, world):
                acc = acc + want.float()
            want = acc.half()
        assert torch.equal(y.view(torch.int16), want.view(torch.int16)), "back-to-back peer allreduce differs"
        if rank == 0:
            print("PEER_OK worst %.2e, large-message allreduce bit-exact" % worst)
    dist.barrier()
    if rank == 0:
        print("TP_OK")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
