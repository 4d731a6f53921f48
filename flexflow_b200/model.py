"""FFModel builders + executor for the attention hot path.

Keeps the reference's builder surface for this path —
  FFModel::dense, inc_multihead_self_attention, spec_inc_multihead_self_attention,
  inc_multihead_self_attention_verify, inc_multiquery_self_attention,
  spec_inc_multiquery_self_attention, inc_multiquery_self_attention_verify
  (include/flexflow/model.h:738-836; layer construction src/ops/inc_multihead_self_attention.cc:56-133)
— and the rule that inserts an AllReduce after `*.o_proj` (and `down_proj`) denses when
tensor_parallelism_degree > 1 (src/runtime/model.cc:3421-3445).  Everything the reference delegates to
Legion (graph search, mapping, futures) is replaced by "run the layers in creation order on the current
CUDA stream of this rank": one process per GPU, the TP shard is this process's rank.
"""
import ctypes as C
from dataclasses import dataclass, field

import torch

from . import batch_config as B
from . import ops


@dataclass
class FFConfig:
    """The FFConfig / RequestManager limits the path reads (src/runtime/model.cc:4425-4570,
    inference/incr_decoding/incr_decoding.cc:97-107)."""
    data_type: torch.dtype = torch.float16
    tensor_parallelism_degree: int = 1
    tp_rank: int = 0
    quantization_type: int = B.DT_NONE
    max_requests_per_batch: int = 8
    max_tokens_per_batch: int = 128
    max_sequence_length: int = 256
    max_spec_tree_token_num: int = B.MAX_SPEC_TREE_TOKEN_NUM
    device: str = "cuda:0"


@dataclass
class Tensor:
    dim: int                      # innermost (hidden) dimension; the outer one is the token dimension
    owner: object = None
    buf: torch.Tensor = None


@dataclass
class Layer:
    op_type: str
    name: str
    inputs: list
    output: Tensor
    attrs: dict = field(default_factory=dict)
    meta: object = None
    weight: torch.Tensor = None
    bias: torch.Tensor = None


class FFModel:
    def __init__(self, ffconfig, comm=None):
        self.config = ffconfig
        self.layers = []
        self.handler = None
        self.comm = comm
        self._compiled = False

    # ---- builders -------------------------------------------------------------------------------
    def create_tensor(self, dim):
        return Tensor(dim)

    def _add(self, op_type, name, inputs, out_dim, **attrs):
        t = Tensor(out_dim)
        lyr = Layer(op_type, name or "%s_%d" % (op_type, len(self.layers)), inputs, t, attrs)
        t.owner = lyr
        self.layers.append(lyr)
        return t

    def dense(self, input, out_dim, use_bias=False, name=None):
        """FFModel::dense (model.h): weight [out_dim][in_dim].  Under TP, `*.qkv_proj`-style denses are
        column-parallel (out_dim is the LOCAL width) and `*.o_proj` denses row-parallel (input is the
        local slice), as the loader shards them (src/runtime/file_loader.cc:253-303,350-354)."""
        return self._add("linear", name, [input], out_dim, use_bias=use_bias)

    def _attn(self, op_type, input, embed_dim, num_q_heads, num_kv_heads, kdim, vdim, dropout, add_zero_attn,
              data_type, kernel_initializer, rotary_embedding_meta, scaling_query, scaling_factor, qk_prod_scaling,
              position_bias, name):
        tp = self.config.tensor_parallelism_degree
        assert num_q_heads % tp == 0, "num_q_heads must divide by the TP degree (inc_…cc:516)"
        head_dim = kdim or embed_dim // num_q_heads
        local_q = num_q_heads // tp
        local_kv = max(1, -(-num_kv_heads // tp))  # ceil, inc_…cc:517-519
        assert input.dim == 3 * local_q * head_dim, "attention input must be the packed [Q|K|V] projection (SURVEY F1)"
        return self._add(op_type, name, [input], local_q * head_dim, num_q_heads=local_q, num_kv_heads=local_kv,
                         global_num_q_heads=num_q_heads, head_dim=head_dim,
                         rotary_embedding_meta=rotary_embedding_meta or ops.RotaryEmbeddingMeta(),
                         scaling_query=scaling_query, scaling_factor=scaling_factor,
                         qk_prod_scaling=qk_prod_scaling, position_bias=position_bias)

    def inc_multiquery_self_attention(self, input, embed_dim, num_q_heads, num_kv_heads, kdim=0, vdim=0,
                                      dropout=0.0, add_zero_attn=False, data_type=B.DT_NONE,
                                      kernel_initializer=None, rotary_embedding_meta=None, scaling_query=False,
                                      scaling_factor=1.0, qk_prod_scaling=True, position_bias=False, name=None):
        return self._attn("inc_attn", input, embed_dim, num_q_heads, num_kv_heads, kdim, vdim, dropout, add_zero_attn,
                          data_type, kernel_initializer, rotary_embedding_meta, scaling_query, scaling_factor,
                          qk_prod_scaling, position_bias, name)

    def spec_inc_multiquery_self_attention(self, input, embed_dim, num_q_heads, num_kv_heads, kdim=0, vdim=0,
                                           dropout=0.0, add_zero_attn=False, data_type=B.DT_NONE,
                                           kernel_initializer=None, rotary_embedding_meta=None, scaling_query=False,
                                           scaling_factor=1.0, qk_prod_scaling=True, position_bias=False, name=None):
        return self._attn("spec_attn", input, embed_dim, num_q_heads, num_kv_heads, kdim, vdim, dropout,
                          add_zero_attn, data_type, kernel_initializer, rotary_embedding_meta, scaling_query,
                          scaling_factor, qk_prod_scaling, position_bias, name)

    def inc_multiquery_self_attention_verify(self, input, embed_dim, num_q_heads, num_kv_heads, kdim=0, vdim=0,
                                             dropout=0.0, add_zero_attn=False, data_type=B.DT_NONE,
                                             kernel_initializer=None, rotary_embedding_meta=None,
                                             scaling_query=False, scaling_factor=1.0, qk_prod_scaling=True,
                                             position_bias=False, name=None):
        return self._attn("tree_attn", input, embed_dim, num_q_heads, num_kv_heads, kdim, vdim, dropout,
                          add_zero_attn, data_type, kernel_initializer, rotary_embedding_meta, scaling_query,
                          scaling_factor, qk_prod_scaling, position_bias, name)

    # multihead == multiquery with num_kv_heads = num_heads (src/ops/inc_multihead_self_attention.cc:56-83)
    def inc_multihead_self_attention(self, input, embed_dim, num_heads, *a, **k):
        return self.inc_multiquery_self_attention(input, embed_dim, num_heads, num_heads, *a, **k)

    def spec_inc_multihead_self_attention(self, input, embed_dim, num_heads, *a, **k):
        return self.spec_inc_multiquery_self_attention(input, embed_dim, num_heads, num_heads, *a, **k)

    def inc_multihead_self_attention_verify(self, input, embed_dim, num_heads, *a, **k):
        return self.inc_multiquery_self_attention_verify(input, embed_dim, num_heads, num_heads, *a, **k)

    # ---- compile + run --------------------------------------------------------------------------
    def compile_inference(self):
        """FFModel::compile_inference (src/runtime/inference_manager.cc:635-777): create op Metas, insert
        AllReduce after o_proj / down_proj denses under TP, allocate activation buffers once."""
        cfg = self.config
        self.handler = ops.FFHandler(cfg.device)
        out = []
        for lyr in self.layers:
            out.append(lyr)
            if (cfg.tensor_parallelism_degree > 1 and lyr.op_type == "linear"
                    and (lyr.name.endswith("o_proj") or lyr.name.endswith("down_proj"))):
                t = Tensor(lyr.output.dim)
                ar = Layer("allreduce", lyr.name + ".allreduce", [lyr.output], t)
                t.owner = ar
                # rewire consumers of the dense output to the all-reduced tensor
                for other in self.layers:
                    other.inputs = [t if (i is lyr.output and other is not ar) else i for i in other.inputs]
                out.append(ar)
        self.layers = out
        attn_meta = {"inc_attn": ops.IncMultiHeadSelfAttentionMeta, "tree_attn": ops.TreeIncMultiHeadSelfAttentionMeta,
                     "spec_attn": ops.SpecIncMultiHeadSelfAttentionMeta}
        with torch.cuda.device(cfg.device):
            for lyr in self.layers:
                lyr.output.buf = torch.zeros(cfg.max_tokens_per_batch, lyr.output.dim, dtype=cfg.data_type,
                                             device=cfg.device)
                if lyr.op_type == "linear":
                    lyr.meta = ops.LinearMeta(self.handler, lyr.inputs[0].dim, lyr.output.dim, cfg.data_type,
                                              cfg.quantization_type, cfg.max_tokens_per_batch)
                elif lyr.op_type in attn_meta:
                    a = lyr.attrs
                    lyr.meta = attn_meta[lyr.op_type](
                        self.handler, a["num_q_heads"], a["num_kv_heads"], a["head_dim"], cfg.data_type,
                        a["rotary_embedding_meta"], a["scaling_query"], a["scaling_factor"], a["qk_prod_scaling"],
                        a["position_bias"], cfg.max_requests_per_batch, cfg.max_tokens_per_batch,
                        cfg.max_sequence_length, cfg.max_spec_tree_token_num, a["global_num_q_heads"])
                elif lyr.op_type == "allreduce":
                    assert self.comm is not None, "tensor parallelism needs a Communicator"
                    lyr.meta = ops.AllReduceMeta(self.handler, self.comm)
        self._compiled = True

    def get_layer(self, name):
        for lyr in self.layers:
            if lyr.name == name:
                return lyr
        raise KeyError(name)

    def set_weight(self, name, weight, bias=None):
        lyr = self.get_layer(name)
        lyr.weight, lyr.bias = weight.contiguous(), bias

    def inference(self, bc, input_tensor, hidden, upload=True):
        """InferenceManager::inference (src/runtime/inference_manager.cc:408-468): upload the batch
        descriptor once, then run every op's inference_kernel_wrapper in order.  `hidden` is the
        [num_tokens][dim] activation feeding `input_tensor`.  Returns the last layer's output rows."""
        assert self._compiled
        T = bc.num_active_tokens()
        if upload:
            self.handler.load_batch_config(bc)
        input_tensor.buf = hidden
        shard = self.config.tp_rank
        fused_into = None  # the AllReduce layer whose work the preceding dense already did
        for li, lyr in enumerate(self.layers):
            x = lyr.inputs[0].buf
            y = lyr.output.buf
            if lyr.op_type == "linear":
                nxt = self.layers[li + 1] if li + 1 < len(self.layers) else None
                if nxt is not None and nxt.op_type == "allreduce" and nxt.inputs[0] is lyr.output:
                    # Linear -> AllReduce as one op (FusedOp, src/ops/fused.cu:448-495): result lands in the
                    # AllReduce layer's output
                    ops.Linear.inference_allreduce_wrapper(lyr.meta, nxt.meta, bc, x, nxt.output.buf, lyr.weight,
                                                           lyr.bias, lyr.inputs[0].dim, lyr.output.dim, T)
                    fused_into = nxt
                    continue
                ops.Linear.inference_kernel_wrapper(lyr.meta, bc, x, y, lyr.weight, lyr.bias, lyr.inputs[0].dim,
                                                    lyr.output.dim, T)
            elif lyr.op_type == "inc_attn":
                ops.IncMultiHeadSelfAttention.inference_kernel_wrapper(lyr.meta, bc, shard, x, y)
            elif lyr.op_type == "tree_attn":
                ops.TreeIncMultiHeadSelfAttention.inference_kernel_wrapper(lyr.meta, bc, shard, x, y)
            elif lyr.op_type == "spec_attn":
                ops.SpecIncMultiHeadSelfAttention.inference_kernel_wrapper(lyr.meta, bc, shard, x, y)
            elif lyr.op_type == "allreduce":
                if lyr is fused_into:
                    fused_into = None
                    continue
                ops.AllReduce.inference_kernel_wrapper(lyr.meta, bc, x[:T], y[:T])
        return self.layers[-1].output.buf[:T]
