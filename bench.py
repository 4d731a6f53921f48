#!/usr/bin/env python
"""bench.py — decode tokens/s of the FlexFlow Serve attention hot path on LLaMA-2-7B shapes.

Workload (BASELINE.json configs[1]): LLaMA-2-7B fp16 incremental decoding, batch 16, 2048-token
context.  One "step" = one decode step of the hot path over all 32 layers:
    qkv_proj dense (M=16) -> IncMultiHeadSelfAttention (RoPE, KV append, decode attention) -> o_proj dense
    [-> AllReduce when TP > 1]
(the MLP / norms / sampling are outside the path).  N GPUs = tensor parallelism N: heads are
sharded, total work is fixed ("strong" scaling).  Weights are random-init, KV caches synthetic.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
os.environ["NCCL_DEBUG"] = "WARN"  # keep stdout to the one JSON line (NCCL prints its version banner there)
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

HIDDEN, NQ, NKV, DH, LAYERS = 4096, 32, 32, 128, 32
BATCH, CTX = 16, 2048
METRIC = "decode tok/s + attn HBM GB/s %peak, LLaMA-2-7B tree-verify, TP=1/2/4/8"
WORKLOAD = "LLaMA-2-7B fp16 incremental decoding, batch 16, 2048-token context (hot path: qkv_proj+attention+o_proj x32 layers)"


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        super().__init__(daemon=True)
        self.gpu, self.rows, self._stop_evt, self.proc = gpu_index, [], threading.Event(), None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([x.strip() for x in line.split(",")])
                if self._stop_evt.is_set():
                    break
        except Exception:
            pass

    def stop(self):
        self._stop_evt.set()
        if self.proc:
            self.proc.terminate()
        sm = sorted(int(float(r[1])) for r in self.rows if len(r) >= 9 and r[1].replace(".", "").isdigit())
        mx = [int(float(r[2])) for r in self.rows if len(r) >= 9 and r[2].replace(".", "").isdigit()]
        reasons = set()
        for r in self.rows:
            if len(r) >= 9:
                for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(name)
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx[0] if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def decode_batch():
    import ffb_testlib as TL
    return TL.make_inc_batch([dict(slot=s, first_depth=CTX - 1, ntok=1) for s in range(BATCH)])


def attn_algorithmic_bytes(n_q_local, n_kv_local):
    """K+V read once + QKV in + out + KV append, per layer per GPU."""
    elt = 2
    kv = 2 * BATCH * CTX * n_kv_local * DH * elt
    qkv_in = BATCH * (n_q_local + 2 * n_kv_local) * DH * elt
    out = BATCH * n_q_local * DH * elt
    append = 2 * BATCH * n_kv_local * DH * elt
    return kv + qkv_in + out + append


def ncu_traffic_bytes(csv_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the committed
    `ncu --set full` capture of this workload (profiles/, same shapes: the traffic does not depend on the run);
    None when the capture is not there."""
    import csv
    path = os.path.join(ROOT, "profiles", csv_name)
    if not os.path.exists(path):
        return None
    try:
        rows = list(csv.reader(open(path)))
        head, units, first = rows[0], rows[1], rows[2]
        mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        tot = 0.0
        for name in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
            i = head.index(name)
            tot += float(first[i].replace(",", "")) * mult[units[i]]
        return tot
    except Exception:  # noqa: BLE001
        return None


def attn_kernel_ms(step_fn, attn_metas, reps):
    """Average duration of the dominant attention kernel: CUDA events recorded around it inside the C ABI
    (ffb_attn_set_profiling; on the stream the kernel is launched on), eager launches.  The GPU is kept ~1 ms behind
    the host so that the events bracket the kernel and not the host's launch latency.  (Events captured into a CUDA
    graph cannot be read back with cudaEventElapsedTime, so the graph replay itself is only timed as a whole.)"""
    import torch
    for m in attn_metas:
        m.set_profiling(True)
    kms = []
    for _ in range(reps):
        torch.cuda._sleep(2_000_000)
        step_fn()
        torch.cuda.synchronize()
        kms.extend(m.last_kernel_ms() for m in attn_metas)
    for m in attn_metas:
        m.set_profiling(False)
    return sum(kms) / len(kms), "events around eager launches, GPU kept behind the host"


def run_ours(args):
    import numpy as np
    import torch
    import torch.distributed as dist

    from flexflow_b200 import batch_config as B
    from flexflow_b200 import ops
    from flexflow_b200.model import FFConfig, FFModel

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus:
        raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (args.gpus, world))
    torch.cuda.set_device(local_rank)
    dev = "cuda:%d" % local_rank
    comm = None
    if world > 1:
        # NCCL writes its version banner to STDOUT at NCCL_DEBUG=VERSION (the box's default); the contract is one JSON
        # line there.  An explicit INFO / TRACE request is left alone.
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=torch.device(dev))

        def bcast(b):
            obj = [b]
            dist.broadcast_object_list(obj, src=0)
            return obj[0]
        comm = ops.Communicator(rank, world, bcast)
    tp = world
    nq_l, nkv_l = NQ // tp, max(1, NKV // tp)
    H_l = nq_l * DH

    cfg = FFConfig(data_type=torch.float16, tensor_parallelism_degree=tp, tp_rank=rank, max_requests_per_batch=BATCH,
                   max_tokens_per_batch=BATCH, max_sequence_length=CTX, device=dev)
    model = FFModel(cfg, comm)
    t_in = model.create_tensor(HIDDEN)
    t = t_in
    rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=10000.0)
    for i in range(LAYERS):  # inference/models/llama.cc:94-196 (attention part of the block)
        qkv = model.dense(t, 3 * H_l, name="layers.%d.self_attn.qkv_proj" % i)
        a = model.inc_multiquery_self_attention(qkv, HIDDEN, NQ, NKV, rotary_embedding_meta=rope,
                                                name="layers.%d.self_attn" % i)
        t = model.dense(a, HIDDEN, name="layers.%d.self_attn.o_proj" % i)
    model.compile_inference()
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    kfill = torch.randn(CTX - 1, nkv_l, DH, generator=g, device=dev, dtype=torch.float16)
    vfill = torch.randn(CTX - 1, nkv_l, DH, generator=g, device=dev, dtype=torch.float16)
    attn_metas = []
    for lyr in model.layers:
        if lyr.op_type == "linear":
            in_dim, out_dim = lyr.inputs[0].dim, lyr.output.dim
            # fan-in of the UNSHARDED layer, so the all-reduced o_proj output has the same scale at every TP degree
            std = (1.0 / HIDDEN) ** 0.5 * (8.0 if lyr.name.endswith("o_proj") else 1.0)
            lyr.weight = (torch.randn(out_dim, in_dim, generator=g, device=dev, dtype=torch.float32) * std).half()
        elif lyr.op_type == "inc_attn":
            attn_metas.append(lyr.meta)
            for s in range(BATCH):
                lyr.meta.write_kv(s, 0, kfill, vfill)
    bc = decode_batch()
    hidden_host = torch.randn(BATCH, HIDDEN, dtype=torch.float16).pin_memory()
    hidden = hidden_host.to(dev)
    result_host = torch.empty(BATCH, HIDDEN, dtype=torch.float16).pin_memory()
    torch.cuda.synchronize()

    def step(upload=False):
        return model.inference(bc, t_in, hidden, upload=upload)

    model.handler.load_batch_config(bc)
    for _ in range(max(args.warmup, 3)):
        out = step()
    torch.cuda.synchronize()
    assert torch.isfinite(out.float()).all(), "non-finite hot-path output"

    graph = None
    if not args.no_graph:
        try:
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                out = step()
            graph.replay()
            torch.cuda.synchronize()
        except Exception as e:  # keep measuring eagerly, say so
            print("[bench] CUDA graph capture failed (%s); timing eager launches" % e, file=sys.stderr)
            graph = None
            torch.cuda.synchronize()

    def run_step():
        if graph is not None:
            graph.replay()
        else:
            step()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        if world > 1:
            tt = torch.tensor([ms], device=dev, dtype=torch.float64)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            return float(tt.item())
        return ms

    # ---- timed region 1: device-resident inputs --------------------------------------------------
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        run_step()
    e1.record()
    barrier()
    ms_dev = max_over_ranks(e0.elapsed_time(e1)) / args.steps

    # ---- timed region 2: end to end through the public API with host buffers ----------------------
    barrier()
    e0.record()
    for _ in range(args.steps):
        hidden.copy_(hidden_host, non_blocking=True)      # H2D of this step's inputs (pinned)
        model.handler.load_batch_config(bc)               # H2D of the batch descriptor (pinned staging)
        run_step()
        result_host.copy_(out, non_blocking=True)         # D2H of the step's result
    e1.record()
    barrier()
    ms_e2e = max_over_ranks(e0.elapsed_time(e1)) / args.steps
    clocks = sampler.stop() if sampler else None
    h2d = BATCH * HIDDEN * 2 + B.CombinedBatchConfigMeta.beamTokenInfo.offset
    d2h = BATCH * HIDDEN * 2

    # ---- roofline of the dominant kernel (decode attention), timed live with CUDA events ----------
    k_ms, k_how = attn_kernel_ms(step, attn_metas, min(args.steps, 10))
    # algorithmic bytes of the decode kernel's launch: K+V rows read once + q in + partial/out write
    alg = attn_algorithmic_bytes(nq_l, nkv_l)
    peak, peak_kind = peaks()
    achieved = alg / (k_ms * 1e-3) / 1e9

    line = {
        "metric": METRIC, "value": BATCH / (ms_dev * 1e-3), "unit": "tokens/s", "n_gpus": world,
        "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": ms_dev, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch": BATCH, "context": CTX, "layers": LAYERS, "parallelism": "tp%d" % tp,
                   "l2_policy": "inputs larger than L2 (17 GB of KV + 4.3 GB of weights stream per step)",
                   "cuda_graph": graph is not None},
        "e2e": {"value": BATCH / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e},
        "gpu_launches": args.steps * LAYERS * 5,
        "clocks": clocks,
        "roofline": {"kernel": "decode_attn_kernel<half,128,8>", "bound": "hbm", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": achieved / peak, "peak_source": peak_kind, "traffic": ncu_traffic_bytes("r01_decode_attn_ncu_full_raw.csv") if tp == 1 else None,
                     "kernel_ms": k_ms, "kernel_timing": k_how, "algorithmic_bytes_per_launch": alg,
                     "attn_share_of_step": k_ms * LAYERS / ms_dev},
    }
    if tp == 1 and not args.no_quantized:
        try:
            line["quantized_decode"] = quantized_extra(args, dev, model, step, ms_dev)
        except Exception as e:  # the headline line must still print
            line["quantized_decode"] = {"error": repr(e)}
    if tp == 1:
        try:
            line["block_ops"] = block_ops_extra(dev)
        except Exception as e:  # the headline line must still print
            line["block_ops"] = {"error": repr(e)}
    if not args.no_tree_verify:
        try:
            line["tree_verify"] = tree_verify_extra(args, dev, comm, rank, tp, model, g)
        except Exception as e:  # the headline line must still print
            line["tree_verify"] = {"error": repr(e)}
    if rank == 0:
        if world == 1 and not args.no_cpu_baseline:
            try:
                line["reference_gpu_kernels"] = reference_gpu_kernels(k_ms)
            except Exception as e:  # the headline line must still print
                line["reference_gpu_kernels"] = {"error": repr(e)}
            line["cpu_baseline"] = cpu_baseline(max_seconds=25.0)
        print(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def quantized_extra(args, dev, model, step, ms_fp16):
    """BASELINE.json configs[2]: the headline decode step with the qkv_proj / o_proj weights stored as the
    reference's int8 / int4 blobs ([values | offsets | scales], groups of 32 along out, fp16 group meta) and
    dequantised in registers inside the weight-streaming kernel.  Synthetic blobs: uniform random values with one
    scale / offset so that the weights span [-0.04, 0.04].  The linear metas of the decode model are swapped for
    quantised ones, the step graph is captured again and timed the same way; restored afterwards."""
    import torch

    from flexflow_b200 import batch_config as B
    from flexflow_b200 import ops

    res = {"workload": "same decode step, qkv_proj / o_proj weights as int8 / int4 blobs, tp1", "fp16_ms_per_step": ms_fp16}
    lins = [l for l in model.layers if l.op_type == "linear"]
    saved = [(l.meta, l.weight) for l in lins]
    gq = torch.Generator(device=dev).manual_seed(99)
    try:
        for name, qt, bits in (("int8", B.DT_INT8, 8), ("int4", B.DT_INT4, 4)):
            wbytes = 0
            for l in lins:
                K, N = l.inputs[0].dim, l.output.dim
                nvals = K * N * bits // 8
                meta = torch.empty(K * N // 32, dtype=torch.float16, device=dev)
                blob = torch.cat([torch.randint(0, 256, (nvals,), generator=gq, device=dev, dtype=torch.uint8),
                                  meta.fill_(-0.04).clone().view(torch.uint8),
                                  meta.fill_((2 ** bits - 1) / 0.08).clone().view(torch.uint8)])
                assert blob.numel() == ops.quant_blob_bytes(qt, K, N), "blob layout"
                l.meta = ops.LinearMeta(model.handler, K, N, torch.float16, qt, BATCH)
                l.weight = blob
                wbytes += blob.numel()
            for _ in range(3):
                out = step()
            torch.cuda.synchronize()
            assert torch.isfinite(out.float()).all(), "non-finite quantised output"
            gr = None
            if not args.no_graph:
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(3):
                gr.replay() if gr is not None else step()
            e0.record()
            for _ in range(args.steps):
                gr.replay() if gr is not None else step()
            e1.record()
            torch.cuda.synchronize()
            ms = e0.elapsed_time(e1) / args.steps

            def lin_only():  # the 64 projection launches alone (weights: 32 distinct layers, no L2 reuse)
                for l in lins:
                    ops.Linear.inference_kernel_wrapper(l.meta, None, l.inputs[0].buf, l.output.buf, l.weight, None,
                                                        l.inputs[0].dim, l.output.dim, BATCH)
            lin_ms = None
            try:
                lg = torch.cuda.CUDAGraph()
                lin_only()
                torch.cuda.synchronize()
                with torch.cuda.graph(lg):
                    lin_only()
                lg.replay()
                a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a0.record()
                for _ in range(5):
                    lg.replay()
                a1.record()
                torch.cuda.synchronize()
                lin_ms = a0.elapsed_time(a1) / 5
            except Exception:  # noqa: BLE001
                torch.cuda.synchronize()
            res[name] = {"ms_per_step": ms, "tokens_per_s": BATCH / (ms * 1e-3), "weight_bytes_per_step": wbytes,
                         "speedup_vs_fp16": ms_fp16 / ms}
            if lin_ms:
                peak, _ = peaks()
                res[name].update({"projections_ms_per_step": lin_ms, "weight_stream_gbs": wbytes / (lin_ms * 1e-3) / 1e9,
                                  "weight_stream_frac_of_hbm_peak": wbytes / (lin_ms * 1e-3) / 1e9 / peak})
            del gr
    finally:
        for l, (m, w) in zip(lins, saved):
            l.meta, l.weight = m, w
    return res


def reference_gpu_kernels(ours_decode_ms):
    """north_star: "the reference's own FlexFlow GPU build is timed too, where it compiles for this box".  The whole
    build needs Legion; what compiles from the reference tree on its own - the decode and tree-verify attention
    kernels - is built into oracle/_ref/ (oracle/Makefile target `ref`, in the container that mounts /root/reference)
    and timed here on the same GPU at the same shapes.  A reported baseline beside cpu_baseline, not the product."""
    import numpy as np
    root = os.path.dirname(os.path.abspath(__file__))
    for p in (root, os.path.join(root, "tests")):
        if p not in sys.path:
            sys.path.insert(0, p)
    from tools import bench_attn as BA
    dec = BA.reference_decode_kernel_ms()
    if dec is None:
        return {"unavailable": "oracle/_ref/libref_attn.so not built (make -C oracle ref needs /root/reference)"}
    tree = BA.reference_tree_kernel_ms(np.random.default_rng(0))
    return {"kind": "reference kernels compiled for sm_100a from the reference tree (oracle/_ref), same GPU",
            "decode_attn": {"workload": "configs[1] shapes: 16 requests x 2048 context, 32 heads x 128, per layer",
                            "reference_kernel_us": dec * 1e3, "ours_us": ours_decode_ms * 1e3,
                            "speedup": dec / ours_decode_ms},
            "tree_verify_attn": {"workload": "8 requests x 25 tree tokens x 512 context (the largest case the reference "
                                             "kernel's shared-memory sizing allows), 32 heads x 128, per layer",
                                 "reference_kernel_us": tree * 1e3,
                                 "note": "ours on the same inputs: tools/bench_attn.py --only treeref "
                                         "(profiles/r01_bench_attn.json)"}}


def block_ops_extra(dev):
    """ResidualRMSNorm (input of qkv_proj) and ArgMax (token ids for accept / reject) at the decode
    and tree-verify batch sizes; graph replay of 40 launches over rotating buffers, algorithmic bytes over the
    measured HBM peak (both are latency-bound at 16 rows)."""
    import torch
    from flexflow_b200 import ops
    peak, _ = peaks()
    res = {}

    def timed(fn, nbuf, reps=40):
        for i in range(3):
            fn(i % nbuf)
        torch.cuda.synchronize()
        gr = torch.cuda.CUDAGraph()
        with torch.cuda.graph(gr):
            for i in range(reps):
                fn(i % nbuf)
        gr.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            gr.replay()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / (5 * reps)

    for rows in (BATCH, 512):
        nbuf = 8
        a = [torch.randn(rows, HIDDEN, device=dev).half() for _ in range(nbuf)]
        b = [torch.randn(rows, HIDDEN, device=dev).half() for _ in range(nbuf)]
        w = torch.ones(HIDDEN, device=dev).half()
        r, o = torch.empty_like(a[0]), torch.empty_like(a[0])
        m = ops.ResidualRMSNormMeta(None, 1e-5, HIDDEN)
        ms = timed(lambda i: ops.ResidualRMSNorm.inference_kernel_wrapper(m, None, a[i], b[i], w, r, o), nbuf)
        alg = rows * HIDDEN * 2 * 4 + HIDDEN * 2
        res["residual_rms_norm_%dx%d" % (rows, HIDDEN)] = {"us": ms * 1e3, "algorithmic_bytes": alg,
                                                            "gbs": alg / (ms * 1e-3) / 1e9,
                                                            "frac_of_hbm_peak": alg / (ms * 1e-3) / 1e9 / peak}
        vocab = 32000
        x = [torch.randn(rows, vocab, device=dev).half() for _ in range(nbuf)]
        idx = torch.empty(rows, dtype=torch.int32, device=dev)
        am = ops.ArgMaxMeta(None)
        ms = timed(lambda i: ops.ArgMax.forward_kernel_wrapper(am, None, x[i], idx, None, rows), nbuf)
        alg = rows * vocab * 2 + rows * 4
        if rows == BATCH:  # the SSM's token selection in beam mode: top 3 of each row
            tk = ops.ArgTopKMeta(None, 3, speculative_decoding=True)
            tidx = torch.empty(rows, 3, dtype=torch.int32, device=dev)
            tval = torch.empty(rows, 3, dtype=torch.float32, device=dev)
            ms3 = timed(lambda i: ops.ArgTopK.forward_kernel_wrapper(tk, None, x[i], tval, tidx, rows), nbuf)
            res["arg_topk3_%dx%d" % (rows, vocab)] = {"us": ms3 * 1e3}
        res["argmax_%dx%d" % (rows, vocab)] = {"us": ms * 1e3, "algorithmic_bytes": alg,
                                               "gbs": alg / (ms * 1e-3) / 1e9,
                                               "frac_of_hbm_peak": alg / (ms * 1e-3) / 1e9 / peak}
    return res


def tree_verify_extra(args, dev, comm, rank, tp, decode_model, g):
    """BASELINE.json configs[3] shapes (LLaMA-2-7B LLM, 8 requests x 64 tree tokens, 1024 committed tokens) through
    the same hot path: qkv_proj (tcgen05 GEMM, M=512) -> TreeIncMultiHeadSelfAttention -> o_proj, 32 layers.
    Reported beside the headline decode line: step time, the tcgen05 attention kernel's algorithmic HBM rate and
    the projection GEMMs' tensor throughput (per-op CUDA-event timing, eager launches)."""
    import numpy as np
    import torch

    import ffb_testlib as TL
    from flexflow_b200 import ops
    from flexflow_b200.model import FFConfig, FFModel

    NREQ, NTREE, COMMITTED = 8, 64, 1024
    T = NREQ * NTREE
    nq_l, nkv_l = NQ // tp, max(1, NKV // tp)
    H_l = nq_l * DH
    cfg = FFConfig(data_type=torch.float16, tensor_parallelism_degree=tp, tp_rank=rank, max_requests_per_batch=NREQ,
                   max_tokens_per_batch=T, max_sequence_length=COMMITTED + 16, max_spec_tree_token_num=NTREE, device=dev)
    model = FFModel(cfg, comm)
    t_in = model.create_tensor(HIDDEN)
    t = t_in
    rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=10000.0)
    for i in range(LAYERS):
        qkv = model.dense(t, 3 * H_l, name="layers.%d.self_attn.qkv_proj" % i)
        a = model.inc_multiquery_self_attention_verify(qkv, HIDDEN, NQ, NKV, rotary_embedding_meta=rope,
                                                       name="layers.%d.self_attn" % i)
        t = model.dense(a, HIDDEN, name="layers.%d.self_attn.o_proj" % i)
    model.compile_inference()
    weights = {l.name: l.weight for l in decode_model.layers if l.op_type == "linear"}
    kfill = torch.randn(COMMITTED, nkv_l, DH, generator=g, device=dev, dtype=torch.float16)
    attn_metas = []
    for lyr in model.layers:
        if lyr.op_type == "linear":
            lyr.weight = weights[lyr.name]  # same random-init weights as the decode model
        elif lyr.op_type == "tree_attn":
            attn_metas.append(lyr.meta)
            for s in range(NREQ):
                lyr.meta.write_kv(s, 0, kfill, kfill)
    rng = np.random.default_rng(0)
    bc = TL.make_tree_batch([dict(slot=s, committed_len=COMMITTED, parent=TL.random_tree(rng, NTREE)) for s in range(NREQ)])
    hidden = torch.randn(T, HIDDEN, generator=g, device=dev, dtype=torch.float16)
    model.handler.load_batch_config(bc)

    def step():
        return model.inference(bc, t_in, hidden, upload=False)
    for _ in range(3):
        out = step()
    torch.cuda.synchronize()
    if not torch.isfinite(out.float()).all():
        bad = [(l.name, int((~torch.isfinite(l.output.buf[:T].float())).sum())) for l in model.layers
               if not torch.isfinite(l.output.buf[:T].float()).all()]
        raise AssertionError("non-finite tree-verify output; first bad layers: %s" % bad[:3])
    graph = None
    try:
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            out = step()
        graph.replay()
        torch.cuda.synchronize()
    except Exception as e:
        print("[bench] tree-verify graph capture failed (%s); timing eager launches" % e, file=sys.stderr)
        graph = None
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    steps = max(5, min(args.steps, 20))
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        graph.replay() if graph is not None else step()
    e1.record()
    torch.cuda.synchronize()
    ms_step = e0.elapsed_time(e1) / steps
    # attention kernel (profiling events inside the C ABI) and per-op GEMM times, eager
    k_ms, k_how = attn_kernel_ms(step, attn_metas, 3)
    lin = {"qkv_proj": [], "o_proj": []}
    x_in = torch.randn(T, HIDDEN, generator=g, device=dev, dtype=torch.float16)
    x_attn = torch.randn(T, H_l, generator=g, device=dev, dtype=torch.float16)
    for kind, x in (("qkv_proj", x_in), ("o_proj", x_attn)):
        lyrs = [l for l in model.layers if l.op_type == "linear" and l.name.endswith(kind)]
        def run_all(lyrs=lyrs, x=x):  # 32 back-to-back launches (weights rotate through 4.3 GB: no L2 reuse)
            for lyr in lyrs:
                ops.Linear.inference_kernel_wrapper(lyr.meta, bc, x, lyr.output.buf, lyr.weight, None, x.shape[1],
                                                    lyr.output.dim, T)
        lg = None
        if graph is not None:
            try:
                torch.cuda.synchronize()
                lg = torch.cuda.CUDAGraph()
                with torch.cuda.graph(lg):
                    run_all()
            except Exception:  # noqa: BLE001
                lg = None
        for rep in range(4):
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            if lg is None:
                torch.cuda._sleep(4_000_000)  # ~2 ms cushion: the host (tensor-map encode + launch) stays ahead
            a0.record()
            lg.replay() if lg is not None else run_all()
            a1.record()
            torch.cuda.synchronize()
            if rep:
                lin[kind].append(a0.elapsed_time(a1) / len(lyrs))
    elt = 2
    kv = 2 * NREQ * (COMMITTED + NTREE) * nkv_l * DH * elt
    alg = kv + T * (nq_l + 2 * nkv_l) * DH * elt + T * nq_l * DH * elt + 2 * T * nkv_l * DH * elt
    peak, _ = peaks()
    tf_peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("bf16_tflops", 1621.0) \
        if os.path.exists(os.path.join(ROOT, "MEASURED_PEAKS.json")) else 1621.0
    res = {"workload": "LLaMA-2-7B tree-verify, 8 requests x 64 tree tokens, 1024 committed tokens, tp%d" % tp,
           "tokens_per_step": T, "ms_per_step": ms_step, "tree_tokens_per_s": T / (ms_step * 1e-3),
           "attn_kernel": "attn_tc_kernel<half,128>", "attn_kernel_ms": k_ms, "attn_algorithmic_bytes": alg,
           "attn_gbs": alg / (k_ms * 1e-3) / 1e9, "attn_frac_of_hbm_peak": alg / (k_ms * 1e-3) / 1e9 / peak,
           "attn_kernel_timing": k_how,
           "cuda_graph": graph is not None}
    for kind, (K_, N_) in (("qkv_proj", (HIDDEN, 3 * H_l)), ("o_proj", (H_l, HIDDEN))):
        ms = sorted(lin[kind])[len(lin[kind]) // 2]
        tf = 2.0 * T * K_ * N_ / (ms * 1e-3) / 1e12
        res[kind] = {"M": T, "K": K_, "N": N_, "ms": ms, "tflops": tf, "frac_of_measured_bf16_peak": tf / tf_peak}
    if tp == 1 and graph is not None:
        # what is left of the graph-replayed step after the graph-timed GEMMs: prep (RoPE + KV append) + attention
        # per layer as they run inside the step (no event, no launch latency): an upper bound of the attention kernel
        rest = ms_step / LAYERS - res["qkv_proj"]["ms"] - res["o_proj"]["ms"]
        res["prep_plus_attn_ms_in_graph_step"] = rest
        res["attn_frac_of_hbm_peak_lower_bound_in_graph_step"] = alg / (rest * 1e-3) / 1e9 / peak
        res["attn_dram_traffic_bytes_ncu"] = ncu_traffic_bytes("r01_attn_tc_ncu_full_raw.csv")
    for m in attn_metas:
        m.close()
    return res


class CpuLayer:
    """One transformer layer's hot path on the CPU oracle (the reference's algorithm restated in C++,
    oracle/*.cpp), at the SAME batch/context as the GPU workload.  A full step is LAYERS of these."""

    def __init__(self, threads):
        import numpy as np
        from flexflow_b200 import batch_config as B
        from oracle import oracle as O
        self.np, self.O = np, O
        O.set_threads(threads)
        rng = np.random.default_rng(0)
        self.attn = O.OracleAttn(B.INC_DECODING_MODE, "half", NQ, DH, BATCH, CTX, 0, BATCH)
        blk = O.round_to_dt(rng.standard_normal((CTX, NQ * DH)).astype(np.float32), "half")
        for s in range(BATCH):
            self.attn.kcache[s, :CTX] = blk
            self.attn.vcache[s, :CTX] = blk[::-1]
        self.wqkv = O.round_to_dt((rng.standard_normal((3 * HIDDEN, HIDDEN)) * 0.0156).astype(np.float32), "half")
        self.wo = O.round_to_dt((rng.standard_normal((HIDDEN, HIDDEN)) * 0.0156).astype(np.float32), "half")
        self.x = O.round_to_dt(rng.standard_normal((BATCH, HIDDEN)).astype(np.float32), "half")
        self.bc = decode_batch()

    def run(self):
        O = self.O
        qkv = O.linear(self.x, self.wqkv, None, "half").reshape(BATCH, 3, HIDDEN)
        a = self.attn.inc_forward(self.bc, qkv)
        return O.linear(a, self.wo, None, "half")


def cpu_baseline(max_seconds=25.0):
    threads = os.cpu_count() or 1
    layer = CpuLayer(threads)
    layer.run()
    t0 = time.perf_counter()
    n = 0
    while True:
        layer.run()
        n += 1
        dt = time.perf_counter() - t0
        if dt > max_seconds or n >= 8:
            break
    per_layer = dt / n
    return {"value": BATCH / (per_layer * LAYERS), "unit": "tokens/s", "cores": threads, "kind": "port",
            "sample": "%d x one layer (qkv_proj+attention+o_proj, batch 16, ctx 2048) of 32, extrapolated x32; "
                      "C++ oracle port, %d threads; the reference itself cannot be built here (Legion absent)" % (n, threads),
            "seconds_per_layer": per_layer}


def run_reference(args):
    """--impl reference: the reference's algorithm on the host cores (oracle port; the reference's own
    build needs Legion, which is not in the tree — SURVEY F4).  Each step = one layer sample, scaled x32."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    layer = CpuLayer(threads)
    for _ in range(min(args.warmup, 1)):
        layer.run()
    steps = max(1, min(args.steps, 6))
    t0 = time.perf_counter()
    for _ in range(steps):
        layer.run()
    per_layer = (time.perf_counter() - t0) / steps
    ms_step = per_layer * LAYERS * 1e3
    val = BATCH / (ms_step * 1e-3)
    sample = ("each step = one layer (qkv_proj+attention+o_proj, batch 16, ctx 2048) timed and scaled x32 layers; "
              "%d timed samples; C++ oracle port with %d threads" % (steps, threads))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": min(args.warmup, 1), "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "batch": BATCH, "context": CTX, "layers": LAYERS, "parallelism": "cpu"},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-tree-verify", action="store_true")
    ap.add_argument("--no-quantized", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
