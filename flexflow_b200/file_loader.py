"""Host mirror of the reference's FileDataLoader for the dense layers either side of the attention operator
(include/flexflow/utils/file_loader.h, src/runtime/file_loader.cc) and of its quantisation tool
(inference/utils/compress_llama_weights.py), over the C ABI's weight-file entry points (include/ffb200.h).  numpy arrays
are plain host buffers here; nothing is computed in Python."""
import ctypes as C
import os

import numpy as np

from . import batch_config as B
from . import lib as _lib

_DT = {"half": (B.DT_HALF, np.uint16), "bf16": (B.DT_BF16, np.uint16), "float": (B.DT_FLOAT, np.float32)}


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


class FileDataLoader:
    """FileDataLoader(weight_file_path, num_q_heads, num_kv_heads, hidden_dim, qkv_inner_dim, tensor_parallelism_degree,
    use_full_precision) - file_loader.cc:27-42.  16-bit tensors are returned as uint16 bit patterns."""

    def __init__(self, weights_folder, num_heads, num_kv_heads, hidden_dim, qkv_inner_dim, tensor_parallelism_degree=1,
                 use_full_precision=False, dtype="half"):
        self.weights_folder = weights_folder
        self.num_heads, self.num_kv_heads = num_heads, num_kv_heads
        self.hidden_dim, self.qkv_inner_dim = hidden_dim, qkv_inner_dim
        self.tensor_parallelism_degree = tensor_parallelism_degree
        self.dtype = "float" if use_full_precision else dtype

    def load_attention_weights_to_dense(self, layer_name, load_o_proj, tp_rank=0):
        """load_attention_weights_to_dense_v2 (file_loader.cc:215-343): the qkv_proj ([3 * H_l][hidden]) or o_proj
        ([hidden][H_l]) weight of TP shard `tp_rank`; tp_rank = -1 gives the reference's whole buffer."""
        code, npdt = _DT[self.dtype]
        tp = self.tensor_parallelism_degree
        H = self.num_heads * self.qkv_inner_dim
        H_l = H // tp
        if load_o_proj:
            shape = (self.hidden_dim, H if tp_rank < 0 else H_l)
        else:
            shape = ((tp if tp_rank < 0 else 1) * 3 * H_l, self.hidden_dim)
        out = np.empty(shape, dtype=npdt)
        _lib.check(_lib.load().ffb_load_attention_weights_to_dense(
            self.weights_folder.encode(), layer_name.encode(), code, self.num_heads, self.num_kv_heads, self.hidden_dim,
            self.qkv_inner_dim, tp, tp_rank, int(bool(load_o_proj)), _p(out)), "ffb_load_attention_weights_to_dense")
        return out

    def load_from_file(self, name, count):
        """load_from_file (file_loader.cc:345-377)."""
        code, npdt = _DT[self.dtype]
        out = np.empty(count, dtype=npdt)
        _lib.check(_lib.load().ffb_load_from_file(os.path.join(self.weights_folder, name).encode(), _p(out), count, code),
                   "ffb_load_from_file")
        return out

    def load_from_quantized_file(self, name, in_dim, out_dim, quantization_type):
        """load_from_quantized_file (file_loader.cc:552-669): the blob [values | offsets | scales] as uint8."""
        code, _ = _DT[self.dtype]
        nbytes = _lib.load().ffb_quant_blob_bytes(quantization_type, code, in_dim, out_dim)
        blob = np.empty(nbytes, dtype=np.uint8)
        _lib.check(_lib.load().ffb_load_from_quantized_file(os.path.join(self.weights_folder, name).encode(),
                                                            quantization_type, code, nbytes, _p(blob)),
                   "ffb_load_from_quantized_file")
        return blob


def compress(weight, num_bits, dtype="half"):
    """compress() of inference/utils/compress_llama_weights.py:15-58 with CompressionConfig(num_bits, group_size=32,
    group_dim=0, symmetric=False) on a [out][in] float32 matrix, packed as the blob the kernels read."""
    w = np.ascontiguousarray(weight, dtype=np.float32)
    out_dim, in_dim = w.shape
    qt = B.DT_INT4 if num_bits == 4 else B.DT_INT8
    code, _ = _DT[dtype]
    blob = np.empty(_lib.load().ffb_quant_blob_bytes(qt, code, in_dim, out_dim), dtype=np.uint8)
    _lib.check(_lib.load().ffb_quantize_weights(_p(w), in_dim, out_dim, qt, code, _p(blob)), "ffb_quantize_weights")
    return blob


def save_quantized(path, blob, in_dim, out_dim, num_bits, dtype="half"):
    """The save step the shipped script lacks: `path`, `path`_offset, `path`_scale as the loader expects them."""
    qt = B.DT_INT4 if num_bits == 4 else B.DT_INT8
    blob = np.ascontiguousarray(blob, dtype=np.uint8)
    _lib.check(_lib.load().ffb_save_quantized_file(path.encode(), _p(blob), in_dim, out_dim, qt, _DT[dtype][0]),
               "ffb_save_quantized_file")
