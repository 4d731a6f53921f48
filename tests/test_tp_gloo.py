"""CPU, world_size 2 over gloo: the head-sharding host logic (flexflow_b200/sharding.py) — shard-local oracle
attention + row-parallel o_proj + allreduce(sum) must equal the unsharded result."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HIDDEN, NQ, NKV, DH, T = 256, 4, 2, 64, 6


def _block(rank, tp, wq, wk, wv, wo, hidden):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import ffb_testlib as TL
    from flexflow_b200 import batch_config as B
    from flexflow_b200.sharding import shard_o_proj, shard_qkv_proj
    from oracle import oracle as O
    hq = NQ // tp
    wqkv = shard_qkv_proj(wq, wk, wv, NQ, NKV, DH, tp, rank).numpy()
    wo_s = shard_o_proj(wo, NQ, DH, tp, rank).numpy()
    qkv = O.linear(hidden, wqkv, None, "float").reshape(T, 3, hq * DH)
    attn = O.OracleAttn(B.INC_DECODING_MODE, "float", hq, DH, 2, 32, 0, 16)
    bc = TL.make_inc_batch([dict(slot=0, first_depth=0, ntok=T, prompt=True)])
    a = attn.inc_forward(bc, qkv)
    return O.linear(a, wo_s, None, "float")


def _worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    g = torch.Generator().manual_seed(3)
    wq = torch.randn(NQ * DH, HIDDEN, generator=g) * 0.05
    wk = torch.randn(NKV * DH, HIDDEN, generator=g) * 0.05
    wv = torch.randn(NKV * DH, HIDDEN, generator=g) * 0.05
    wo = torch.randn(HIDDEN, NQ * DH, generator=g) * 0.05
    hidden = torch.randn(T, HIDDEN, generator=g).numpy()
    part = torch.from_numpy(_block(rank, world, wq, wk, wv, wo, hidden))
    dist.all_reduce(part)  # the AllReduce parallel op after o_proj (model.cc:3421-3445)
    if rank == 0:
        full = _block(0, 1, wq, wk, wv, wo, hidden)
        ret["err"] = float(np.abs(part.numpy() - full).max())
    dist.barrier()
    dist.destroy_process_group()


def test_tp2_gloo_equals_unsharded():
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, 29571, ret), nprocs=2, join=True)
    assert ret["err"] < 1e-4, ret["err"]
