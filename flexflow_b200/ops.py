"""Host-side mirror of the reference's operator interface for the attention hot path.

Same names, argument meaning and call order as the reference's Legion-free core:
  IncMultiHeadSelfAttention.inference_kernel_wrapper(m, bc, shard_id, input, output)
      include/flexflow/ops/inc_multihead_self_attention.h:115-119
  TreeIncMultiHeadSelfAttention / SpecIncMultiHeadSelfAttention  (…tree_…h:104-108, …spec_…h:101-106)
  Linear.inference_kernel_wrapper(m, bc, input, output, weight, bias, in_dim, out_dim, batch)
      src/ops/kernels/linear_kernels.cu:187-195
  AllReduce.inference_kernel_wrapper(m, bc, input, output)
      include/flexflow/parallel_ops/kernels/allreduce_kernels.h:28-31
Tensors are torch CUDA tensors used purely as device buffers (the role of GenericTensorAccessorR/W);
all compute happens in libffb200.so.  Errors raise FFBError (the reference asserts/exits).
"""
import ctypes as C
from dataclasses import dataclass

import torch

from . import batch_config as B
from . import lib as L


@dataclass
class RotaryEmbeddingMeta:
    """include/flexflow/inference.h:46-66"""
    apply_rotary_embedding: bool = False
    rope_theta: float = 10000.0
    rope_type: str = "default"
    factor: float = 8.0
    low_freq_factor: float = 1.0
    high_freq_factor: float = 4.0
    original_max_position_embeddings: int = 8192


def _dt_of(t):
    if t.dtype == torch.float16:
        return B.DT_HALF
    if t.dtype == torch.bfloat16:
        return B.DT_BF16
    raise L.FFBError("unsupported tensor dtype %s (half/bf16 only)" % t.dtype)


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda or not t.is_contiguous():
        raise L.FFBError("expected a contiguous CUDA tensor")
    return C.c_void_p(t.data_ptr())


class FFHandler:
    """The slice of FFHandler the path uses (include/flexflow/config.h:91-113): the device-resident
    CombinedBatchConfigMetaStruct plus a pinned staging block for the single-shot upload."""

    def __init__(self, device="cuda:0"):
        self.device = torch.device(device)
        n = C.sizeof(B.CombinedBatchConfigMeta)
        self.batch_config_metadata = torch.zeros(n, dtype=torch.uint8, device=self.device)
        self._staging = torch.zeros(n, dtype=torch.uint8).pin_memory()
        self.ncclComm = None

    def load_batch_config(self, bc):
        """RequestManager::load_batch_config_task (src/runtime/request_manager.cu:83-165)."""
        mode = mode_of(bc)
        lib = L.load()
        L.check(lib.ffb_load_batch_config(C.addressof(bc), mode, self._staging.data_ptr(),
                                          self.batch_config_metadata.data_ptr(), _stream()),
                "ffb_load_batch_config")


def mode_of(bc):
    if isinstance(bc, B.TreeVerifyBatchConfig):
        return B.TREE_VERIFY_MODE
    if isinstance(bc, B.BeamSearchBatchConfig):
        return B.BEAM_SEARCH_MODE
    return B.INC_DECODING_MODE


class IncMultiHeadSelfAttentionMeta:
    """IncMultiHeadSelfAttentionMeta (inc_multihead_self_attention.h:139-202; ctor inc_…cu:1621-1807).
    num_q_heads / num_kv_heads are the per-shard counts (inc_…cc:516-519)."""

    MODE = B.INC_DECODING_MODE

    def __init__(self, handler, num_q_heads, num_kv_heads, head_dim, dtype=torch.float16,
                 rotary_embedding_meta=None, scaling_query=False, scaling_factor=1.0, qk_prod_scaling=True,
                 position_bias=False, max_requests_per_batch=8, max_tokens_per_batch=128,
                 max_sequence_length=256, max_spec_tree_token_num=B.MAX_SPEC_TREE_TOKEN_NUM,
                 global_num_q_heads=None):
        self.handle = handler
        rm = rotary_embedding_meta or RotaryEmbeddingMeta()
        d = L.AttnDesc()
        d.mode = self.MODE
        d.dtype = B.DT_HALF if dtype == torch.float16 else (B.DT_BF16 if dtype == torch.bfloat16 else B.DT_FLOAT)
        d.num_q_heads, d.num_kv_heads, d.head_dim = num_q_heads, num_kv_heads, head_dim
        d.global_num_q_heads = global_num_q_heads or num_q_heads
        d.apply_rotary_embedding = int(rm.apply_rotary_embedding)
        d.rope_theta = rm.rope_theta
        d.rope_llama3 = int(rm.rope_type == "llama3")
        d.rope_factor, d.rope_low_freq_factor, d.rope_high_freq_factor = rm.factor, rm.low_freq_factor, rm.high_freq_factor
        d.rope_original_max_position_embeddings = rm.original_max_position_embeddings
        d.scaling_query, d.scaling_factor = int(scaling_query), scaling_factor
        d.qk_prod_scaling, d.position_bias = int(qk_prod_scaling), int(position_bias)
        d.max_requests, d.max_sequence_length = max_requests_per_batch, max_sequence_length
        d.max_spec_tree_tokens = 0 if self.MODE == B.INC_DECODING_MODE else max_spec_tree_token_num
        d.max_tokens_per_batch = max_tokens_per_batch
        self.desc = d
        self.dtype = dtype
        self.num_q_heads, self.num_kv_heads, self.head_dim = num_q_heads, num_kv_heads, head_dim
        self.hidden_size = num_q_heads * head_dim
        self._h = C.c_void_p()
        with torch.cuda.device(handler.device):
            L.check(L.load().ffb_attn_create(C.byref(d), C.byref(self._h)), "ffb_attn_create")

    def device_bytes(self):
        return L.load().ffb_attn_device_bytes(self._h)

    def set_profiling(self, enable):
        L.check(L.load().ffb_attn_set_profiling(self._h, int(enable)), "ffb_attn_set_profiling")

    def last_kernel_ms(self):
        ms = C.c_float(0)
        L.check(L.load().ffb_attn_last_kernel_ms(self._h, C.byref(ms)), "ffb_attn_last_kernel_ms")
        return ms.value

    def write_kv(self, request, pos0, k, v):
        """Test helper: load K/V rows [n][num_kv_heads][head_dim] into the private cache."""
        L.check(L.load().ffb_attn_write_kv(self._h, request, pos0, k.shape[0], _ptr(k), _ptr(v), _stream()),
                "ffb_attn_write_kv")

    def read_kv(self, request, pos0, n):
        k = torch.empty(n, self.num_kv_heads, self.head_dim, dtype=self.dtype, device=self.handle.device)
        v = torch.empty_like(k)
        L.check(L.load().ffb_attn_read_kv(self._h, request, pos0, n, _ptr(k), _ptr(v), _stream()),
                "ffb_attn_read_kv")
        return k, v

    def close(self):
        if self._h:
            L.load().ffb_attn_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class TreeIncMultiHeadSelfAttentionMeta(IncMultiHeadSelfAttentionMeta):
    """tree_inc_multihead_self_attention.cu:670-718"""
    MODE = B.TREE_VERIFY_MODE


class SpecIncMultiHeadSelfAttentionMeta(IncMultiHeadSelfAttentionMeta):
    """spec_inc_multihead_self_attention.cu:769-818"""
    MODE = B.BEAM_SEARCH_MODE


def _attn_call(fn_name, m, bc, shard_id, input, output):
    if input.dtype != output.dtype:
        raise L.FFBError("input/output dtype mismatch")  # assert(input.data_type == output.data_type)
    fn = getattr(L.load(), fn_name)
    L.check(fn(m._h, C.addressof(bc), m.handle.batch_config_metadata.data_ptr(), shard_id, _ptr(input),
               _ptr(output), _stream()), fn_name)


class IncMultiHeadSelfAttention:
    @staticmethod
    def inference_kernel_wrapper(m, bc, shard_id, input, output):
        _attn_call("ffb_inc_attn_forward", m, bc, shard_id, input, output)


class TreeIncMultiHeadSelfAttention:
    @staticmethod
    def inference_kernel_wrapper(m, bc, shard_id, input, output):
        _attn_call("ffb_tree_attn_forward", m, bc, shard_id, input, output)


class SpecIncMultiHeadSelfAttention:
    @staticmethod
    def inference_kernel_wrapper(m, bc, shard_id, input, output):
        _attn_call("ffb_spec_attn_forward", m, bc, shard_id, input, output)


class LinearMeta:
    """LinearMeta for the forward path (include/flexflow/ops/kernels/linear_kernels.h)."""

    def __init__(self, handler, in_dim, out_dim, dtype=torch.float16, quantization_type=B.DT_NONE,
                 max_tokens_per_batch=128):
        self.handle = handler
        d = L.LinearDesc(B.DT_HALF if dtype == torch.float16 else B.DT_BF16, quantization_type, in_dim, out_dim,
                         max_tokens_per_batch, 0)
        self.desc = d
        self.in_dim, self.out_dim, self.dtype = in_dim, out_dim, dtype
        self._h = C.c_void_p()
        with torch.cuda.device(handler.device):
            L.check(L.load().ffb_linear_create(C.byref(d), C.byref(self._h)), "ffb_linear_create")

    def close(self):
        if self._h:
            L.load().ffb_linear_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Linear:
    @staticmethod
    def forward_kernel_wrapper(m, input, output, weight, bias, in_dim, out_dim, batch_size):
        if in_dim != m.in_dim or out_dim != m.out_dim:
            raise L.FFBError("shape does not match LinearMeta")
        L.check(L.load().ffb_linear_forward(m._h, _ptr(input), _ptr(output), _ptr(weight), _ptr(bias), batch_size,
                                            _stream()), "ffb_linear_forward")

    @staticmethod
    def inference_kernel_wrapper(m, bc, input, output, weight, bias, in_dim, out_dim, batch_size):
        Linear.forward_kernel_wrapper(m, input, output, weight, bias, in_dim, out_dim, batch_size)

    @staticmethod
    def inference_allreduce_wrapper(m, ar_meta, bc, input, output, weight, bias, in_dim, out_dim, batch_size):
        """Linear -> AllReduce run back to back by FusedOp::inference_task (src/ops/fused.cu:448-495): one fused
        kernel over NVLink peer memory for decode batches, otherwise linear + NCCL allreduce (ffb200.h)."""
        if in_dim != m.in_dim or out_dim != m.out_dim:
            raise L.FFBError("shape does not match LinearMeta")
        L.check(L.load().ffb_linear_forward_allreduce(m._h, ar_meta.comm._h, _ptr(input), _ptr(output), _ptr(weight),
                                                      _ptr(bias), batch_size, _stream()), "ffb_linear_forward_allreduce")


class AllReduceMeta:
    """AllReduceMeta: holds the communicator (handle.ncclComm in the reference)."""

    def __init__(self, handler, comm):
        self.handle = handler
        self.comm = comm


class AllReduce:
    @staticmethod
    def inference_kernel_wrapper(m, bc, input, output):
        """allreduce_kernels.cu:56-78: num_elements = bc->num_active_tokens() * hidden_dim."""
        if input.dtype != output.dtype or input.shape != output.shape:
            raise L.FFBError("allreduce input/output mismatch")
        hidden = input.shape[-1]
        count = bc.num_active_tokens() * hidden
        L.check(L.load().ffb_allreduce(m.comm._h, _ptr(input), _ptr(output), count, _dt_of(input), _stream()),
                "ffb_allreduce")


class Communicator:
    """NCCL communicator bring-up (src/runtime/model.cc:584-628): unique id from rank 0, init on all ranks.
    `broadcast_bytes(bytes_or_None) -> bytes` distributes the 128-byte id (e.g. torch.distributed)."""

    def __init__(self, rank, nranks, broadcast_bytes):
        lib = L.load()
        buf = (C.c_char * 128)()
        if rank == 0:
            L.check(lib.ffb_comm_get_unique_id(buf), "ffb_comm_get_unique_id")
        uid = broadcast_bytes(bytes(buf) if rank == 0 else None)
        self._h = C.c_void_p()
        L.check(lib.ffb_comm_init_rank(uid, nranks, rank, C.byref(self._h)), "ffb_comm_init_rank")
        self.rank, self.nranks = rank, nranks
        self.has_peer = bool(lib.ffb_comm_has_peer(self._h))

    def close(self):
        if self._h:
            L.load().ffb_comm_destroy(self._h)
            self._h = C.c_void_p()


def quant_blob_bytes(quant_type, in_dim, out_dim, dtype=torch.float16):
    return L.load().ffb_quant_blob_bytes(quant_type, B.DT_HALF if dtype == torch.float16 else B.DT_BF16, in_dim, out_dim)


def decompress_weights(blob, in_dim, out_dim, quant_type, dtype=torch.float16):
    """decompress_int{4,8}_general_weights (src/ops/kernels/decompress_kernels.cu:27-70) -> [out][in] DT."""
    out = torch.empty(out_dim, in_dim, dtype=dtype, device=blob.device)
    fn = L.load().ffb_dequant_int4 if quant_type == B.DT_INT4 else L.load().ffb_dequant_int8
    L.check(fn(_ptr(blob), _ptr(out), in_dim, out_dim, B.DT_HALF if dtype == torch.float16 else B.DT_BF16, _stream()),
            "ffb_dequant")
    return out


class ResidualRMSNormMeta:
    """ResidualRMSNormMeta / RMSNormMeta (include/flexflow/ops/residual_rms_norm.h, rms_norm.h): eps and in_dim; the
    reference's rms_ptr / norm_ptr scratch does not exist here (the row stays in registers)."""

    def __init__(self, handler, eps, in_dim):
        self.handle, self.eps, self.in_dim = handler, float(eps), int(in_dim)


class ResidualRMSNorm:
    @staticmethod
    def inference_kernel_wrapper(m, bc, input1, input2, weight, residual_output, output):
        """Kernels::ResidualRMSNorm::inference_kernel_wrapper (src/ops/kernels/residual_rms_norm_kernels.cu:205-232)."""
        n = input1.numel() // m.in_dim
        L.check(L.load().ffb_residual_rms_norm(_ptr(input1), _ptr(input2), _ptr(weight), _ptr(residual_output),
                                               _ptr(output), n, m.in_dim, m.eps, _dt_of(input1), _stream()),
                "ffb_residual_rms_norm")


class RMSNorm:
    @staticmethod
    def inference_kernel_wrapper(m, bc, input, weight, output):
        """Kernels::RMSNorm::inference_kernel_wrapper (src/ops/kernels/rms_norm_kernels.cu:148-190)."""
        n = input.numel() // m.in_dim
        L.check(L.load().ffb_residual_rms_norm(_ptr(input), C.c_void_p(0), _ptr(weight), C.c_void_p(0), _ptr(output),
                                               n, m.in_dim, m.eps, _dt_of(input), _stream()), "ffb_residual_rms_norm")


class SoftmaxMeta:
    """SoftmaxMeta (include/flexflow/ops/softmax.h): dim is always the class dimension on this path."""

    def __init__(self, handler, dim=-1):
        self.handle, self.dim = handler, dim


class Softmax:
    @staticmethod
    def inference_kernel_wrapper(m, bc, is_last_op, input, output, output_grad=None):
        """Kernels::Softmax::inference_kernel_wrapper (src/ops/kernels/softmax.cu:121-166): input / output
        [num_tokens][num_classes] DT.  (output_grad, the PEFT copy of the last op, is outside the path.)"""
        num_classes = input.shape[-1]
        n = input.numel() // num_classes
        L.check(L.load().ffb_softmax(_ptr(input), _ptr(output), n, num_classes, _dt_of(input), _stream()), "ffb_softmax")


class ArgMaxMeta:
    """ArgMaxMeta (include/flexflow/ops/argmax.h): beam_search selects the prob / parent outputs."""

    def __init__(self, handler, beam_search=False):
        self.handle, self.beam_search = handler, bool(beam_search)


class ArgMax:
    @staticmethod
    def forward_kernel_wrapper(m, bc, input, indices, parent, batch_size):
        """ArgMax::forward_kernel_wrapper (src/ops/argmax.cu:156-206): input [batch_size][length] DT, indices int32
        [batch_size]; returns the probs tensor in the beam-search case (the reference keeps it in m->probs)."""
        length = input.numel() // batch_size
        if indices.dtype != torch.int32:
            raise L.FFBError("indices must be int32")
        probs = torch.empty(batch_size, dtype=torch.float32, device=input.device) if m.beam_search else None
        L.check(L.load().ffb_argmax(_ptr(input), _ptr(indices), _ptr(probs), _ptr(parent) if m.beam_search else
                                    C.c_void_p(0), batch_size, length, _dt_of(input), _stream()), "ffb_argmax")
        return probs


class ArgTopKMeta:
    """ArgTopKMeta (include/flexflow/ops/arg_topk.h): k, sorted (always best first here), speculative_decoding selects
    the values output."""

    def __init__(self, handler, k, speculative_decoding=False):
        self.handle, self.k, self.speculative_decoding = handler, int(k), bool(speculative_decoding)


class ArgTopK:
    @staticmethod
    def forward_kernel_wrapper(m, bc, input, probs, indices, batch_size):
        """ArgTopK::forward_kernel_wrapper (src/ops/arg_topk.cu:451-541): input [batch_size][length] DT, indices int32
        [batch_size][k], probs float32 [batch_size][k] (written in the speculative-decoding case)."""
        length = input.numel() // batch_size
        if indices.dtype != torch.int32:
            raise L.FFBError("indices must be int32")
        L.check(L.load().ffb_arg_topk(_ptr(input), _ptr(probs) if m.speculative_decoding else C.c_void_p(0), _ptr(indices),
                                      batch_size, length, m.k, _dt_of(input), _stream()), "ffb_arg_topk")
