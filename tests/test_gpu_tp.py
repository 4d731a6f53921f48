"""GPU (>= 2 devices): head-sharded tensor parallelism — qkv_proj column-parallel, attention shard-local,
o_proj row-parallel, NCCL allreduce — must reproduce the unsharded result (src/runtime/model.cc:3421-3445,
src/runtime/file_loader.cc:253-303)."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_tp2_matches_tp1():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(ROOT, "tests", "tp_worker.py")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "TP_OK" in r.stdout


def test_tp2_fused_oproj_allreduce_over_nvlink():
    """o_proj + AllReduce as one kernel over NVLink peer memory (decode batches) vs linear + NCCL and vs TP=1."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, TP_DIMS="2048,16,16,128", TP_EXPECT_PEER="1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29534",
                        os.path.join(ROOT, "tests", "tp_worker.py")], capture_output=True, text=True, timeout=300, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "TP_OK" in r.stdout and "PEER_OK" in r.stdout
