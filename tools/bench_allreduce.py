"""AllReduce micro-benchmark at N GPUs (torchrun): ffb_allreduce (peer-memory kernel / NCCL fallback) vs torch NCCL.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/bench_allreduce.py [--mb 4]"""
import ctypes as C
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from flexflow_b200 import batch_config as B  # noqa: E402
from flexflow_b200 import lib as L  # noqa: E402
from flexflow_b200 import ops  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    dev = "cuda:%d" % int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(dev)
    dist.init_process_group("nccl", device_id=torch.device(dev))

    def bcast(b):
        obj = [b]
        dist.broadcast_object_list(obj, src=0)
        return obj[0]
    comm = ops.Communicator(rank, world, bcast)
    sizes = [float(sys.argv[sys.argv.index("--mb") + 1])] if "--mb" in sys.argv else [0.25, 1, 2, 4, 8]
    iters = 20
    for mb in sizes:
        n = int(mb * (1 << 20)) // 2
        bufs = [torch.randn(n, device=dev).half() for _ in range(4)]
        outs = [torch.empty(n, dtype=torch.float16, device=dev) for _ in range(4)]
        st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

        def ours(i):
            L.check(L.load().ffb_allreduce(comm._h, C.c_void_p(bufs[i % 4].data_ptr()), C.c_void_p(outs[i % 4].data_ptr()), n,
                                           B.DT_HALF, st), "ffb_allreduce")

        def nccl(i):
            dist.all_reduce(outs[i % 4])
        res = {}
        for name, fn in (("ours", ours), ("nccl", nccl)):
            for i in range(3):
                fn(i)
            torch.cuda.synchronize()
            dist.barrier()
            torch.cuda.synchronize()
            ts = []
            for rep in range(3):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(3_000_000)
                e0.record()
                for i in range(iters):
                    fn(i)
                e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1) / iters * 1e3)
            t = torch.tensor([sorted(ts)[1]], device=dev, dtype=torch.float64)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[name] = float(t.item())
        if rank == 0:
            print({"mb": mb, "world": world, "has_peer": bool(comm.has_peer), "ours_us": res["ours"], "nccl_us": res["nccl"],
                   "ours_busbw_gbs": 2 * (world - 1) / world * mb * 1.048576 / res["ours"] * 1e3})
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
