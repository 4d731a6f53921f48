#!/usr/bin/env python
"""bench.py — the FlexFlow Serve attention hot path on LLaMA-2-7B shapes, headline = the config BASELINE.json's metric
is quoted on: tree-verify (SpecInfer's verification step).

Headline workload (BASELINE.json configs[3]): LLaMA-2-7B LLM, 8 requests x 64 speculated tree tokens, 1024 committed
tokens per request, fp16.  One "step" = one verification step of the hot path over all 32 layers:
    qkv_proj dense (M = 512) -> TreeIncMultiHeadSelfAttention (RoPE, KV append, tree-masked attention) -> o_proj dense
    [-> AllReduce when TP > 1]
(norms / MLP / sampling are outside the path).  N GPUs = tensor parallelism N: heads are sharded, the total
work is fixed ("strong" scaling).  Weights are random-init, KV caches and trees synthetic.
`value`  = tree tokens verified per second, inputs resident in HBM (CUDA-graph replay of the step).
`e2e`    = the same through the public API (FFModel.inference) with HOST buffers, EAGER launches (what a per-operator
           Legion integration sees): every step copies the 4 MiB of hidden states H2D from pinned memory, uploads the
           TreeVerifyBatchConfig, runs the 32 layers and copies the 4 MiB result D2H.  `e2e_graph` replays a CUDA graph
           instead of launching eagerly (valid only while the batch descriptor does not change).
Extras on the same line: configs[1] incremental decode (b16 x 2048), configs[2] int8 / int4 weights, the row operators,
per-layer time split, TP parity of both paths at N > 1, configs[4] (70B shard shapes) at N = 8, the reference's own
kernels on the same GPU, and the CPU baselines (HuggingFace/PyTorch path on the host cores).

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
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

HIDDEN, NQ, NKV, DH, LAYERS = 4096, 32, 32, 128, 32
BATCH, CTX = 16, 2048                    # configs[1]
NREQ, NTREE, COMMITTED = 8, 64, 1024     # configs[3]
TREE_T = NREQ * NTREE
METRIC = "decode tok/s + attn HBM GB/s %peak, LLaMA-2-7B tree-verify, TP=1/2/4/8"
WORKLOAD = ("LLaMA-2-7B fp16 tree-verify (SpecInfer verification step), 8 requests x 64 tree tokens, 1024 committed tokens "
            "(hot path: qkv_proj + TreeIncMultiHeadSelfAttention + o_proj x 32 layers)")
DECODE_WORKLOAD = "LLaMA-2-7B fp16 incremental decoding, batch 16, 2048-token context (same hot path, 32 layers)"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("bf16_tflops", 1657.2)), "measured"
    return 6650.0, 1657.2, "fallback"


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


def attn_algorithmic_bytes(kv_rows_total, T, n_q_local, n_kv_local, elt=2):
    """K+V read once + QKV in + out + KV append, per layer per GPU.  kv_rows_total = sum over requests of
    the rows attended (committed + new)."""
    return (2 * kv_rows_total * n_kv_local * DH * elt + T * (n_q_local + 2 * n_kv_local) * DH * elt
            + T * n_q_local * DH * elt + 2 * T * n_kv_local * DH * elt)


class Dist:
    """One process per GPU: rank / world, the ffb communicator (NCCL + peer mapping), barrier and max-over-ranks."""

    def __init__(self, gpus):
        import torch
        import torch.distributed as dist
        from flexflow_b200 import ops
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if self.world != gpus:
            raise SystemExit("--gpus %d but WORLD_SIZE=%d (launch with torch.distributed.run)" % (gpus, self.world))
        torch.cuda.set_device(self.local_rank)
        self.dev = "cuda:%d" % self.local_rank
        self.comm, self.dist, self.torch = None, dist, torch
        if self.world > 1:
            # NCCL_DEBUG is whatever the launcher set (the driver reads NCCL's own rank lines); stdout may carry NCCL's
            # banner first, this program's ONE JSON line comes last.
            dist.init_process_group("nccl", device_id=torch.device(self.dev))

            def bcast(b):
                obj = [b]
                dist.broadcast_object_list(obj, src=0)
                return obj[0]
            self.comm = ops.Communicator(self.rank, self.world, bcast)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_ms(self, ms):
        if self.world > 1:
            tt = self.torch.tensor([ms], device=self.dev, dtype=self.torch.float64)
            self.dist.all_reduce(tt, op=self.dist.ReduceOp.MAX)
            return float(tt.item())
        return ms

    def close(self):
        if self.world > 1:
            self.dist.barrier()
            self.dist.destroy_process_group()


class Workload:
    """An N-layer hot-path model (qkv_proj -> attention -> o_proj [-> allreduce]) + one synthetic batch + host buffers."""

    def __init__(self, D, kind, layers=LAYERS, weights=None, hidden_dim=HIDDEN, nq=NQ, nkv=NKV, nreq=None,
                 ntree=None, committed=None, name="layers"):
        import numpy as np
        import torch
        import ffb_testlib as TL
        from flexflow_b200 import ops
        from flexflow_b200.model import FFConfig, FFModel
        self.D, self.kind, self.nlayers, self.hidden_dim = D, kind, layers, hidden_dim
        tp, rank, dev = D.world, D.rank, D.dev
        self.nq, self.nkv = nq, nkv
        self.nq_l, self.nkv_l = nq // tp, max(1, nkv // tp)
        self.H_l = self.nq_l * DH
        g = torch.Generator(device=dev).manual_seed(1234 + rank)
        if kind == "decode":
            self.T, self.nreq, self.rows = BATCH, BATCH, CTX
            cfg = FFConfig(data_type=torch.float16, tensor_parallelism_degree=tp, tp_rank=rank, max_requests_per_batch=BATCH,
                           max_tokens_per_batch=BATCH, max_sequence_length=CTX, device=dev)
        else:
            self.nreq, ntree, committed = nreq or NREQ, ntree or NTREE, committed or COMMITTED
            self.T, self.rows = self.nreq * ntree, committed + ntree
            cfg = FFConfig(data_type=torch.float16, tensor_parallelism_degree=tp, tp_rank=rank,
                           max_requests_per_batch=self.nreq, max_tokens_per_batch=self.T,
                           max_sequence_length=committed + 16, max_spec_tree_token_num=ntree, device=dev)
        self.model = model = FFModel(cfg, D.comm)
        self.t_in = model.create_tensor(hidden_dim)
        t = self.t_in
        rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=10000.0)
        for i in range(layers):  # inference/models/llama.cc:94-196 (attention part of the block)
            qkv = model.dense(t, 3 * self.H_l, name="%s.%d.self_attn.qkv_proj" % (name, i))
            if kind == "decode":
                a = model.inc_multiquery_self_attention(qkv, hidden_dim, nq, nkv, kdim=DH, rotary_embedding_meta=rope,
                                                        name="%s.%d.self_attn" % (name, i))
            else:
                a = model.inc_multiquery_self_attention_verify(qkv, hidden_dim, nq, nkv, kdim=DH, rotary_embedding_meta=rope,
                                                               name="%s.%d.self_attn" % (name, i))
            t = model.dense(a, hidden_dim, name="%s.%d.self_attn.o_proj" % (name, i))
        model.compile_inference()
        fill_rows = CTX - 1 if kind == "decode" else committed
        self.kfill = torch.randn(fill_rows, self.nkv_l, DH, generator=g, device=dev, dtype=torch.float16)
        self.vfill = torch.randn(fill_rows, self.nkv_l, DH, generator=g, device=dev, dtype=torch.float16)
        self.attn_metas, self.lins = [], []
        for lyr in model.layers:
            if lyr.op_type == "linear":
                self.lins.append(lyr)
                if weights is not None and lyr.name in weights:
                    lyr.weight = weights[lyr.name]
                else:
                    in_dim, out_dim = lyr.inputs[0].dim, lyr.output.dim
                    # fan-in of the UNSHARDED layer, so the all-reduced o_proj output has the same scale at every TP degree
                    std = (1.0 / hidden_dim) ** 0.5 * (8.0 if lyr.name.endswith("o_proj") else 1.0)
                    lyr.weight = (torch.randn(out_dim, in_dim, generator=g, device=dev, dtype=torch.float32) * std).half()
            elif lyr.op_type in ("inc_attn", "tree_attn"):
                self.attn_metas.append(lyr.meta)
                for s in range(self.nreq):
                    lyr.meta.write_kv(s, 0, self.kfill, self.vfill)
        if kind == "decode":
            self.bc = TL.make_inc_batch([dict(slot=s, first_depth=CTX - 1, ntok=1) for s in range(BATCH)])
        else:
            rng = np.random.default_rng(0)
            self.bc = TL.make_tree_batch([dict(slot=s, committed_len=committed, parent=TL.random_tree(rng, ntree))
                                          for s in range(self.nreq)])
        cpu_g = torch.Generator().manual_seed(4321)  # the same activations on every rank (replicated input, model.cc:3124)
        self.hidden_host = torch.randn(self.T, hidden_dim, generator=cpu_g).half().pin_memory()
        self.hidden = self.hidden_host.to(dev)
        self.result_host = torch.empty(self.T, hidden_dim, dtype=torch.float16).pin_memory()
        self.graph, self.out = None, None
        torch.cuda.synchronize()

    def weights(self):
        return {l.name: l.weight for l in self.lins}

    def step(self, upload=False):
        self.out = self.model.inference(self.bc, self.t_in, self.hidden, upload=upload)
        return self.out

    def warm(self, n, try_graph=True):
        import torch
        self.model.handler.load_batch_config(self.bc)
        for _ in range(n):
            out = self.step()
        torch.cuda.synchronize()
        if not torch.isfinite(out.float()).all():
            bad = [(l.name, int((~torch.isfinite(l.output.buf[:self.T].float())).sum())) for l in self.model.layers
                   if not torch.isfinite(l.output.buf[:self.T].float()).all()]
            raise AssertionError("non-finite %s output; first bad layers: %s" % (self.kind, bad[:3]))
        if try_graph:
            try:
                self.graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self.graph):
                    self.step()
                self.graph.replay()
                torch.cuda.synchronize()
            except Exception as e:  # keep measuring eagerly, say so
                print("[bench] CUDA graph capture of the %s step failed (%s); timing eager launches" % (self.kind, e), file=sys.stderr)
                self.graph = None
                torch.cuda.synchronize()

    def run_resident(self):
        if self.graph is not None:
            self.graph.replay()
        else:
            self.step()

    def time_resident(self, steps):
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.D.barrier()
        e0.record()
        for _ in range(steps):
            self.run_resident()
        e1.record()
        self.D.barrier()
        return self.D.max_ms(e0.elapsed_time(e1)) / steps

    def time_e2e(self, steps, graph):
        """Host buffers, host<->device copies inside the timed region, through FFModel.inference (eager) or a replay."""
        import torch
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.D.barrier()
        e0.record()
        for _ in range(steps):
            self.hidden.copy_(self.hidden_host, non_blocking=True)             # H2D of this step's inputs (pinned)
            if graph:
                self.model.handler.load_batch_config(self.bc)                  # H2D of the batch descriptor
                self.graph.replay()
            else:
                self.step(upload=True)                                         # descriptor upload + all layers, eager
            self.result_host.copy_(self.out, non_blocking=True)                # D2H of the step's result
        e1.record()
        self.D.barrier()
        return self.D.max_ms(e0.elapsed_time(e1)) / steps

    def e2e_bytes(self):
        from flexflow_b200 import batch_config as B
        desc = B.CombinedBatchConfigMeta.beamTokenInfo.offset if self.kind == "decode" else C.sizeof(B.CombinedBatchConfigMeta)
        return self.T * self.hidden_dim * 2 + desc, self.T * self.hidden_dim * 2

    def attn_kernel_ms(self, reps):
        """Average duration of the dominant attention kernel: CUDA events recorded around it inside the C ABI
        (ffb_attn_set_profiling; on the stream the kernel is launched on), eager launches.  The GPU is kept ~1 ms
        behind the host so that the events bracket the kernel and not the host's launch latency.  (Events captured
        into a CUDA graph cannot be read back, so the graph replay itself is only timed as a whole.)"""
        import torch
        for m in self.attn_metas:
            m.set_profiling(True)
        kms = []
        for _ in range(reps):
            torch.cuda._sleep(2_000_000)
            self.step()
            torch.cuda.synchronize()
            kms.extend(m.last_kernel_ms() for m in self.attn_metas)
        for m in self.attn_metas:
            m.set_profiling(False)
        kms.sort()
        return sum(kms) / len(kms), kms[len(kms) // 2], "events around eager launches, GPU kept behind the host"

    def attn_kernel_run_ms(self):
        """The same kernel timed as a RUN: ffb_attn_set_profiling(2) makes every attention call launch only its
        dominant kernel (no RoPE / KV-append kernel, no dependent launch), so one event pair brackets the kernels of
        all layers back to back - each layer streams its own cache (no L2 reuse) - and the quotient does not carry
        the ~5 us a single-launch event bracket adds (event record -> launch -> completion -> event record)."""
        import torch
        from flexflow_b200 import ops
        attn_op = ops.IncMultiHeadSelfAttention if self.kind == "decode" else ops.TreeIncMultiHeadSelfAttention
        attn_layers = [l for l in self.model.layers if l.op_type in ("inc_attn", "tree_attn")]
        self.step()  # a full step leaves q / descriptors / appended rows in every handle
        torch.cuda.synchronize()
        x_qkv = attn_layers[0].inputs[0].buf
        for m in self.attn_metas:
            m.set_profiling(2)
        times = []
        try:
            for rep in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                torch.cuda._sleep(4_000_000)
                e0.record()
                for l in attn_layers:
                    attn_op.inference_kernel_wrapper(l.meta, self.bc, self.D.rank, x_qkv, l.output.buf)
                e1.record()
                torch.cuda.synchronize()
                if rep:
                    times.append(e0.elapsed_time(e1) / len(attn_layers))
        finally:
            for m in self.attn_metas:
                m.set_profiling(False)
        return sorted(times)[len(times) // 2]

    def layer_split_us(self):
        """Per-layer time of the three parts of the chain, each timed alone as a graph of its launches (weights and
        caches rotate through all layers: no L2 reuse): qkv_proj, attention operator (prep + attention kernels),
        o_proj (+ AllReduce under TP)."""
        import torch
        from flexflow_b200 import ops
        model, bc, T = self.model, self.bc, self.T
        g = torch.Generator(device=self.D.dev).manual_seed(77)
        x_in = torch.randn(T, self.hidden_dim, generator=g, device=self.D.dev, dtype=torch.float16)
        x_attn = torch.randn(T, self.H_l, generator=g, device=self.D.dev, dtype=torch.float16)
        x_qkv = torch.randn(T, 3 * self.H_l, generator=g, device=self.D.dev, dtype=torch.float16)
        layers = model.layers
        attn_op = ops.IncMultiHeadSelfAttention if self.kind == "decode" else ops.TreeIncMultiHeadSelfAttention

        def qkv_all():
            for l in layers:
                if l.op_type == "linear" and l.name.endswith("qkv_proj"):
                    ops.Linear.inference_kernel_wrapper(l.meta, bc, x_in, l.output.buf, l.weight, None, l.inputs[0].dim, l.output.dim, T)

        def attn_all():
            for l in layers:
                if l.op_type in ("inc_attn", "tree_attn"):
                    attn_op.inference_kernel_wrapper(l.meta, bc, self.D.rank, x_qkv, l.output.buf)

        def o_all():
            for i, l in enumerate(layers):
                if l.op_type == "linear" and l.name.endswith("o_proj"):
                    nxt = layers[i + 1] if i + 1 < len(layers) else None
                    if nxt is not None and nxt.op_type == "allreduce":
                        ops.Linear.inference_allreduce_wrapper(l.meta, nxt.meta, bc, x_attn, nxt.output.buf, l.weight, None,
                                                               l.inputs[0].dim, l.output.dim, T)
                    else:
                        ops.Linear.inference_kernel_wrapper(l.meta, bc, x_attn, l.output.buf, l.weight, None, l.inputs[0].dim, l.output.dim, T)
        res = {}
        for name, fn in (("qkv_proj", qkv_all), ("attention_op", attn_all), ("o_proj_allreduce" if self.D.world > 1 else "o_proj", o_all)):
            fn()
            torch.cuda.synchronize()
            gr = None
            try:
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    fn()
            except Exception:  # noqa: BLE001
                gr = None
                torch.cuda.synchronize()
            times = []
            for rep in range(4):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                self.D.barrier()
                if gr is None:
                    torch.cuda._sleep(4_000_000)
                e0.record()
                gr.replay() if gr is not None else fn()
                e1.record()
                torch.cuda.synchronize()
                if rep:
                    times.append(e0.elapsed_time(e1) / self.nlayers * 1e3)
            res[name] = self.D.max_ms(sorted(times)[len(times) // 2])
        return res

    def close(self):
        self.graph = None
        for lyr in self.model.layers:
            if lyr.meta is not None and hasattr(lyr.meta, "close"):
                lyr.meta.close()


def tp_parity(W):
    """Driver-visible TP parity (N > 1): layer 0 of the sharded model (qkv_proj -> attention -> o_proj + AllReduce, the
    path the timed step takes: fused peer push for decode batches, GEMM + ncclAllReduce for tree-verify) against the
    SAME layer recomputed unsharded on rank 0 from the gathered weight and KV shards, plus a bit comparison of every
    rank's all-reduced output."""
    import torch
    import torch.distributed as dist
    from flexflow_b200 import ops
    from flexflow_b200.model import FFConfig, FFModel
    D, tp = W.D, W.D.world
    model, bc, T = W.model, W.bc, W.T
    layers = model.layers
    qkv_l, attn_l, o_l, ar_l = layers[0], layers[1], layers[2], layers[3]
    assert ar_l.op_type == "allreduce"
    model.handler.load_batch_config(bc)
    ops.Linear.inference_kernel_wrapper(qkv_l.meta, bc, W.hidden, qkv_l.output.buf, qkv_l.weight, None, W.hidden_dim, 3 * W.H_l, T)
    aop = ops.IncMultiHeadSelfAttention if W.kind == "decode" else ops.TreeIncMultiHeadSelfAttention
    aop.inference_kernel_wrapper(attn_l.meta, bc, D.rank, qkv_l.output.buf, attn_l.output.buf)
    ops.Linear.inference_allreduce_wrapper(o_l.meta, ar_l.meta, bc, attn_l.output.buf, ar_l.output.buf, o_l.weight, None,
                                           W.H_l, W.hidden_dim, T)
    torch.cuda.synchronize()
    mine = ar_l.output.buf[:T].clone()

    def gather(t):
        parts = [torch.empty_like(t) for _ in range(tp)]
        dist.all_gather(parts, t.contiguous())
        return parts
    outs = gather(mine)
    identical = all(torch.equal(outs[0].view(torch.int16), o.view(torch.int16)) for o in outs)
    wq = gather(qkv_l.weight)      # shard r: [Q_r | K_r | V_r] rows (file_loader.cc:253-303)
    wo = gather(o_l.weight)        # shard r: columns of its heads (file_loader.cc:350-354)
    kf, vf = gather(W.kfill), gather(W.vfill)
    res = {"ranks_bit_identical": bool(identical),
           "path": ("o_proj + ncclAllReduce" if not D.comm.has_peer else
                    ("fused o_proj + NVLink peer-push allreduce (one kernel)" if W.kind == "decode" else
                     "o_proj GEMM + two-shot peer-memory allreduce kernel (reduce-scatter / all-gather by NVLink push)"))}
    err = torch.zeros(1, device=D.dev, dtype=torch.float64)
    if D.rank == 0:
        H_l = W.H_l
        w_full = torch.cat([torch.cat([w[j * H_l:(j + 1) * H_l] for w in wq], 0) for j in range(3)], 0).contiguous()
        wo_full = torch.cat(wo, 1).contiguous()
        cfg = FFConfig(data_type=torch.float16, tensor_parallelism_degree=1, tp_rank=0, max_requests_per_batch=W.nreq,
                       max_tokens_per_batch=T, max_sequence_length=model.config.max_sequence_length,
                       max_spec_tree_token_num=model.config.max_spec_tree_token_num, device=D.dev)
        one = FFModel(cfg, None)
        t_in = one.create_tensor(W.hidden_dim)
        rope = ops.RotaryEmbeddingMeta(apply_rotary_embedding=True, rope_theta=10000.0)
        q = one.dense(t_in, 3 * W.nq * DH, name="l0.qkv_proj")
        # distinct kv heads of the unsharded model = every shard's, in rank order (same q heads per kv head as a shard)
        a = (one.inc_multiquery_self_attention if W.kind == "decode" else one.inc_multiquery_self_attention_verify)(
            q, W.hidden_dim, W.nq, W.nkv_l * tp, kdim=DH, rotary_embedding_meta=rope, name="l0.attn")
        one.dense(a, W.hidden_dim, name="l0.o_proj")
        one.compile_inference()
        one.layers[0].weight, one.layers[2].weight = w_full, wo_full
        kfull, vfull = torch.cat(kf, 1).contiguous(), torch.cat(vf, 1).contiguous()
        for s in range(W.nreq):
            one.layers[1].meta.write_kv(s, 0, kfull, vfull)
        want = one.inference(bc, t_in, W.hidden, upload=True).float()
        torch.cuda.synchronize()
        err[0] = float((mine.float() - want).abs().max() / want.abs().max().clamp_min(1e-12))
        for l in one.layers:
            if l.meta is not None and hasattr(l.meta, "close"):
                l.meta.close()
    dist.broadcast(err, src=0)
    res["rel_err_vs_unsharded"] = float(err.item())
    if res["rel_err_vs_unsharded"] > 1e-2 or not identical:
        raise AssertionError("TP parity failed for the %s path: %s" % (W.kind, res))
    return res


def quantized_extra(args, W, ms_fp16):
    """BASELINE.json configs[2]: the decode step with the qkv_proj / o_proj weights stored as the reference's int8 /
    int4 blobs ([values | offsets | scales], groups of 32 along out, fp16 group meta) and dequantised on chip inside
    the weight-streaming kernel.  Synthetic blobs: uniform random values with one scale / offset so that the weights
    span [-0.04, 0.04].  The linear metas of the decode model are swapped for quantised ones, the step graph is
    captured again and timed the same way; restored afterwards."""
    import torch
    from flexflow_b200 import batch_config as B
    from flexflow_b200 import ops
    dev, model = W.D.dev, W.model
    res = {"workload": "configs[1] decode step with qkv_proj / o_proj weights as int8 / int4 blobs, tp1", "fp16_ms_per_step": ms_fp16}
    lins = W.lins
    saved = [(l.meta, l.weight) for l in lins]
    gq = torch.Generator(device=dev).manual_seed(99)
    hbm_peak = measured_peaks()[0]
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
                out = W.step()
            torch.cuda.synchronize()
            assert torch.isfinite(out.float()).all(), "non-finite quantised output"
            gr = None
            if not args.no_graph:
                gr = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gr):
                    W.step()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            for _ in range(3):
                gr.replay() if gr is not None else W.step()
            e0.record()
            for _ in range(args.steps):
                gr.replay() if gr is not None else W.step()
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
                res[name].update({"projections_ms_per_step": lin_ms, "weight_stream_gbs": wbytes / (lin_ms * 1e-3) / 1e9,
                                  "weight_stream_frac_of_hbm_peak": wbytes / (lin_ms * 1e-3) / 1e9 / hbm_peak})
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
                                 "note": "ours on the same inputs: tools/bench_attn.py --only treeref (profiles/)"}}


def block_ops_extra(dev):
    """ResidualRMSNorm (input of qkv_proj) and ArgMax (token ids for accept / reject) at the decode
    and tree-verify batch sizes; graph replay of 40 launches over rotating buffers, algorithmic bytes over the
    measured HBM peak (both are latency-bound at 16 rows)."""
    import torch
    from flexflow_b200 import ops
    peak = measured_peaks()[0]
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


def run_ours(args):
    import torch

    D = Dist(args.gpus)
    tp, rank, dev = D.world, D.rank, D.dev
    warmup = max(args.warmup, 3)
    hbm_peak, tf_peak, peak_kind = measured_peaks()

    # ---- headline: configs[3] tree-verify ------------------------------------------------------------------------
    tv = Workload(D, "tree")
    tv.warm(warmup, try_graph=not args.no_graph)
    sampler = ClockSampler(D.local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
        time.sleep(0.3)
    ms_dev = tv.time_resident(args.steps)
    for _ in range(2):
        tv.step(upload=True)
    ms_e2e = tv.time_e2e(args.steps, graph=False)
    ms_e2e_graph = tv.time_e2e(args.steps, graph=True) if tv.graph is not None else None
    clocks = sampler.stop() if sampler else None
    h2d, d2h = tv.e2e_bytes()
    k1_ms, k1_med, k1_how = tv.attn_kernel_ms(min(args.steps, 6))
    k_ms = tv.attn_kernel_run_ms()
    alg = attn_algorithmic_bytes(NREQ * (COMMITTED + NTREE), TREE_T, tv.nq_l, tv.nkv_l)
    achieved = alg / (k_ms * 1e-3) / 1e9
    split_tv = tv.layer_split_us()
    launches_per_layer = 4  # qkv GEMM, prep (RoPE + KV append), attention, o_proj GEMM (+ NCCL's own kernel under TP)
    line = {
        "metric": METRIC, "value": TREE_T / (ms_dev * 1e-3), "unit": "tokens/s", "n_gpus": tp,
        "steps": args.steps, "warmup": warmup, "ms_per_step": ms_dev, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f16", "data": "synthetic",
        "config": {"workload": WORKLOAD, "requests": NREQ, "tree_tokens_per_request": NTREE, "committed_tokens": COMMITTED,
                   "tokens_per_step": TREE_T, "layers": LAYERS, "parallelism": "tp%d" % tp,
                   "tokens": "tree tokens verified per second; accepted tokens/s = value x (accepted per request and step) / 64",
                   "l2_policy": "inputs larger than L2 (4.6 GB of KV + 4.3 GB of weights stream per step)",
                   "cuda_graph": tv.graph is not None},
        "e2e": {"value": TREE_T / (ms_e2e * 1e-3), "unit": "tokens/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                "ms_per_step": ms_e2e, "launch": "eager (FFModel.inference: descriptor upload + 32 x 3 operator calls per step)"},
        "e2e_graph": None if ms_e2e_graph is None else {
            "value": TREE_T / (ms_e2e_graph * 1e-3), "unit": "tokens/s", "ms_per_step": ms_e2e_graph,
            "note": "CUDA-graph replay; the tile schedule is baked at capture, so it is valid only for an unchanged batch descriptor"},
        "gpu_launches": args.steps * LAYERS * launches_per_layer,
        "clocks": clocks,
        "roofline": {"kernel": "attn_tc_kernel<half,128>", "bound": "hbm", "achieved": achieved, "peak": hbm_peak,
                     "unit": "GB/s", "frac": achieved / hbm_peak, "peak_source": peak_kind,
                     "traffic": ncu_traffic_bytes("r02_attn_tc_ncu_full_raw.csv") if tp == 1 else None,
                     "kernel_ms": k_ms,
                     "kernel_timing": "one CUDA-event pair around the %d layers' attention kernels launched back to back "
                                      "(ffb_attn_set_profiling(2): that kernel only, each layer's own cache), / %d" % (LAYERS, LAYERS),
                     "single_launch_bracket": {"kernel_ms": k1_ms, "kernel_ms_median": k1_med, "how": k1_how,
                                               "frac": alg / (k1_ms * 1e-3) / 1e9 / hbm_peak,
                                               "note": "event pair inside the C ABI around ONE launch inside the step: "
                                                       "adds the record -> launch -> completion -> record latency (~5 us)"},
                     "algorithmic_bytes_per_launch": alg, "attn_share_of_step": k_ms * LAYERS / ms_dev,
                     "frac_of_nominal_8TBs": achieved / 8000.0},
        "projections": {}, "layer_split_us": {"tree_verify": split_tv},
    }
    for kind, K_, N_ in (("qkv_proj", HIDDEN, 3 * tv.H_l), ("o_proj", tv.H_l, HIDDEN)):
        key = kind if kind in split_tv else "o_proj_allreduce"
        us = split_tv[key]
        tf = 2.0 * TREE_T * K_ * N_ / (us * 1e-6) / 1e12
        line["projections"][kind] = {"M": TREE_T, "K": K_, "N": N_, "us": us, "tflops": tf, "frac_of_measured_bf16_peak": tf / tf_peak,
                                     "includes_allreduce": key != kind}
    if tp > 1:
        line["tp_parity"] = {"tree_verify": tp_parity(tv)}
    weights = tv.weights()

    # ---- configs[1]: incremental decode --------------------------------------------------------------------------
    if not args.no_decode:
        try:
            dec = Workload(D, "decode", weights=weights)
            dec.warm(warmup, try_graph=not args.no_graph)
            d_ms = dec.time_resident(args.steps)
            for _ in range(2):
                dec.step(upload=True)
            d_e2e = dec.time_e2e(args.steps, graph=False)
            d_e2e_g = dec.time_e2e(args.steps, graph=True) if dec.graph is not None else None
            dk1_ms, dk_med, _ = dec.attn_kernel_ms(min(args.steps, 6))
            dk_ms = dec.attn_kernel_run_ms()
            d_alg = attn_algorithmic_bytes(BATCH * CTX, BATCH, dec.nq_l, dec.nkv_l)
            d_h2d, d_d2h = dec.e2e_bytes()
            line["decode"] = {
                "workload": DECODE_WORKLOAD + ", tp%d" % tp, "tokens_per_step": BATCH, "ms_per_step": d_ms,
                "tokens_per_s": BATCH / (d_ms * 1e-3), "cuda_graph": dec.graph is not None,
                "e2e": {"value": BATCH / (d_e2e * 1e-3), "unit": "tokens/s", "ms_per_step": d_e2e, "launch": "eager",
                        "h2d_bytes_per_step": d_h2d, "d2h_bytes_per_step": d_d2h},
                "e2e_graph": None if d_e2e_g is None else {"value": BATCH / (d_e2e_g * 1e-3), "ms_per_step": d_e2e_g},
                "roofline": {"kernel": "decode_attn_kernel<half,128,8>", "bound": "hbm", "achieved": d_alg / (dk_ms * 1e-3) / 1e9,
                             "peak": hbm_peak, "unit": "GB/s", "frac": d_alg / (dk_ms * 1e-3) / 1e9 / hbm_peak,
                             "kernel_ms": dk_ms, "single_launch_bracket_kernel_ms": dk1_ms, "single_launch_bracket_median_ms": dk_med,
                             "algorithmic_bytes_per_launch": d_alg,
                             "traffic": ncu_traffic_bytes("r02_decode_attn_ncu_full_raw.csv") if tp == 1 else None,
                             "attn_share_of_step": dk_ms * LAYERS / d_ms}}
            line["layer_split_us"]["decode"] = dec.layer_split_us()
            if tp > 1:
                line["tp_parity"]["decode"] = tp_parity(dec)
            if tp == 1 and not args.no_quantized:
                try:
                    line["quantized_decode"] = quantized_extra(args, dec, d_ms)
                except Exception as e:  # the headline line must still print
                    line["quantized_decode"] = {"error": repr(e)}
            if rank == 0 and tp == 1 and not args.no_cpu_baseline:
                try:
                    line["reference_gpu_kernels"] = reference_gpu_kernels(dk_ms)
                except Exception as e:  # noqa: BLE001
                    line["reference_gpu_kernels"] = {"error": repr(e)}
            dec.close()
            del dec
        except AssertionError:
            raise
        except Exception as e:  # the headline line must still print
            line["decode"] = {"error": repr(e)}
    if tp == 1:
        try:
            line["block_ops"] = block_ops_extra(dev)
        except Exception as e:  # noqa: BLE001
            line["block_ops"] = {"error": repr(e)}
    tv.close()
    del tv
    torch.cuda.empty_cache()

    # ---- configs[4]: LLaMA-2-70B shard shapes at TP = 8 ------------------------------------------------------------
    if tp == 8 and not args.no_70b:
        try:
            L70 = 80
            big = Workload(D, "tree", layers=L70, hidden_dim=8192, nq=64, nkv=8, nreq=32, ntree=32, committed=4096, name="l70b")
            big.warm(3, try_graph=not args.no_graph)
            b_ms = big.time_resident(max(3, min(args.steps, 10)))
            bk_ms, bk_med, _ = big.attn_kernel_ms(2)
            b_alg = attn_algorithmic_bytes(32 * (4096 + 32), 1024, big.nq_l, big.nkv_l)
            line["llama70b_tree_verify_tp8"] = {
                "workload": "BASELINE configs[4]: LLaMA-2-70B fp16 tree-verify, 32 requests x 32 tree tokens (T = 1024 = "
                            "MAX_NUM_TOKENS, SURVEY F10), 4096 committed, 80 layers, 8 q heads + 1 kv head per GPU",
                "ms_per_step": b_ms, "tree_tokens_per_s": 1024 / (b_ms * 1e-3), "cuda_graph": big.graph is not None,
                "attn_kernel_ms": bk_ms, "attn_algorithmic_bytes": b_alg, "attn_gbs": b_alg / (bk_ms * 1e-3) / 1e9,
                "attn_frac_of_hbm_peak": b_alg / (bk_ms * 1e-3) / 1e9 / hbm_peak,
                "layer_split_us": big.layer_split_us()}
            try:
                line["llama70b_tree_verify_tp8"]["tp_parity"] = tp_parity(big)
            except AssertionError:
                raise
            except Exception as e:  # noqa: BLE001
                line["llama70b_tree_verify_tp8"]["tp_parity"] = {"error": repr(e)}
            big.close()
        except AssertionError:
            raise
        except Exception as e:  # noqa: BLE001
            line["llama70b_tree_verify_tp8"] = {"error": repr(e)}

    if rank == 0:
        if tp == 1 and not args.no_cpu_baseline:
            try:
                line["cpu_baseline"] = hf_cpu_baseline(max_seconds=25.0)
            except Exception as e:  # noqa: BLE001
                line["cpu_baseline"] = {"error": repr(e)}
        sys.stdout.flush()
        print(json.dumps(line), flush=True)
    D.close()


# ---- CPU baselines: the HuggingFace/PyTorch path the reference's alignment tests use ---------------------------------
class HFTreeVerifyLayerCPU:
    """The hot path of ONE LLaMA-2-7B decoder layer (q/k/v/o projections, RoPE, attention over 1024 cached + 64 new
    positions with the tree mask) as HuggingFace's LlamaAttention module computes it in fp32 on the host cores
    (tests/inference/huggingface_inference.py:57-101 loads the same modelling code through AutoModelForCausalLM; the
    reference's alignment tests hook exactly this module, tests/peft/hf_utils.py:117-168).  A step of the workload is
    32 such layers; the same module is applied 32 times (its 268 MB of weights do not fit any cache level)."""

    def __init__(self, threads):
        import numpy as np
        import torch
        from transformers import LlamaConfig
        from transformers.cache_utils import DynamicCache
        from transformers.models.llama.modeling_llama import LlamaAttention, LlamaRotaryEmbedding
        import ffb_testlib as TL
        torch.set_num_threads(threads)
        self.torch, self.DynamicCache = torch, DynamicCache
        cfg = LlamaConfig(hidden_size=HIDDEN, intermediate_size=11008, num_hidden_layers=1, num_attention_heads=NQ,
                          num_key_value_heads=NKV, vocab_size=32000, max_position_embeddings=4096, rope_theta=10000.0,
                          attention_bias=False)
        cfg._attn_implementation = "sdpa"
        torch.manual_seed(0)
        self.attn = LlamaAttention(cfg, layer_idx=0).eval().float()
        rot = LlamaRotaryEmbedding(cfg)
        rng = np.random.default_rng(0)
        trees = [TL.random_tree(rng, NTREE) for _ in range(NREQ)]
        depth = np.stack([np.array(TL.tree_depths(t)) for t in trees])
        pos = torch.from_numpy(COMMITTED + depth).long()
        self.x = torch.randn(NREQ, NTREE, HIDDEN)
        self.pos_emb = rot(self.x, pos)
        self.past_k = torch.randn(NREQ, NKV, COMMITTED, DH)
        self.past_v = torch.randn(NREQ, NKV, COMMITTED, DH)
        mask = torch.zeros(NREQ, 1, NTREE, COMMITTED + NTREE)
        for r in range(NREQ):
            anc = TL.ancestor_mask(trees[r])
            for k in range(NTREE):
                for q in range(NTREE):
                    if not (anc[k] >> q) & 1:
                        mask[r, 0, q, COMMITTED + k] = float("-inf")
        self.mask = mask

    def layer(self):
        torch = self.torch
        cache = self.DynamicCache()
        cache.update(self.past_k, self.past_v, 0)
        with torch.no_grad():
            out, _ = self.attn(self.x, position_embeddings=self.pos_emb, attention_mask=self.mask, past_key_values=cache)
        return out

    def step(self):
        for _ in range(LAYERS):
            out = self.layer()
        return out


def hf_configs0_cpu(threads):
    """BASELINE configs[0] in full: LLaMA-160M architecture (random init, seed 0), fp32, one 16-token prompt, 64 new
    tokens, HuggingFace generate(do_sample=False) on the host cores (huggingface_inference.py:86-101).  Best of 3."""
    import torch
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    import make_hf_decode_golden as G
    torch.set_num_threads(threads)
    model = G.build_model(0)
    ids = G.prompt_ids(0)
    best = None
    for _ in range(3):
        t0 = time.perf_counter()
        with torch.no_grad():
            out = model.generate(ids, max_new_tokens=64, min_new_tokens=64, do_sample=False, num_beams=1, pad_token_id=0)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    n_new = int(out.shape[1] - ids.shape[1])
    return {"workload": "configs[0]: LLaMA-160M arch fp32, 1 prompt x 16 tokens, %d new tokens, HF generate greedy" % n_new,
            "seconds": best, "tokens_per_s": n_new / best, "timing": "wall clock, best of 3, whole generate() call"}


def hf_cpu_baseline(max_seconds=25.0):
    import torch
    threads = os.cpu_count() or 1
    lay = HFTreeVerifyLayerCPU(threads)
    lay.layer()
    t0 = time.perf_counter()
    n = 0
    while True:
        lay.layer()
        n += 1
        dt = time.perf_counter() - t0
        if dt > max_seconds * 0.6 or n >= 64:
            break
    per_layer = dt / n
    res = {"value": TREE_T / (per_layer * LAYERS), "unit": "tokens/s", "cores": threads, "kind": "hf",
           "torch_threads": torch.get_num_threads(), "dtype": "f32",
           "sample": "%d x one LLaMA-2-7B decoder layer's hot path (HuggingFace LlamaAttention: q/k/v/o projections, RoPE, "
                     "SDPA over 1024 cached + 64 tree tokens x 8 requests, tree mask) of the 32 in a step, extrapolated x32; "
                     "`bench.py --impl reference` times whole 32-layer steps" % n,
           "seconds_per_layer": per_layer, "extrapolated": True}
    try:
        res["configs0"] = hf_configs0_cpu(threads)
    except Exception as e:  # noqa: BLE001
        res["configs0"] = {"error": repr(e)}
    return res


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path = the HuggingFace/PyTorch alignment path
    (north_star; the FlexFlow C++/CUDA build needs Legion, which is not in the tree - SURVEY F4) on the host cores, on
    the headline workload.  A step = a whole 32-layer verification step (no extrapolation); --steps / --warmup are
    honoured unless the run would exceed ~5 minutes, in which case the timed steps are cut and the line says so."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    threads = os.cpu_count() or 1
    lay = HFTreeVerifyLayerCPU(threads)
    budget = float(os.environ.get("FFB_REFERENCE_BUDGET_S", "280"))
    t_begin = time.perf_counter()
    warm = 0
    per_step = None
    want_warm = max(args.warmup, 1)
    for _ in range(want_warm):
        t0 = time.perf_counter()
        lay.step()
        per_step = time.perf_counter() - t0
        warm += 1
        if (time.perf_counter() - t_begin) + per_step * (args.steps + (want_warm - warm)) > budget:
            break  # the remaining warm-up steps would eat the budget
    left = budget - (time.perf_counter() - t_begin)
    steps = max(1, min(args.steps, int(left / max(per_step, 1e-9))))
    t0 = time.perf_counter()
    for _ in range(steps):
        lay.step()
    ms_step = (time.perf_counter() - t0) / steps * 1e3
    val = TREE_T / (ms_step * 1e-3)
    sample = ("each step = one whole verification step: HuggingFace LlamaAttention (q/k/v/o projections, RoPE, SDPA with the "
              "tree mask over 1024 cached + 64 tree tokens x 8 requests) applied 32 times, fp32, %d torch threads; %d timed "
              "steps after %d warm-up%s" % (torch.get_num_threads(), steps, warm,
                                            "" if steps == args.steps else " (cut from --steps %d to stay inside %.0f s)" % (args.steps, budget)))
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "tokens/s", "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "requests": NREQ, "tree_tokens_per_request": NTREE, "committed_tokens": COMMITTED,
                   "tokens_per_step": TREE_T, "layers": LAYERS, "parallelism": "cpu"},
        "cpu_baseline": {"value": val, "unit": "tokens/s", "cores": threads, "kind": "hf", "sample": sample,
                         "torch_threads": torch.get_num_threads()},
        "e2e": {"value": val, "unit": "tokens/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-decode", action="store_true")
    ap.add_argument("--no-quantized", action="store_true")
    ap.add_argument("--no-70b", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
