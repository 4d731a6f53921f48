"""ctypes binding of libffb200.so (the C ABI in include/ffb200.h).

There is no CPU fallback: if the CUDA extension is missing or a call fails, an exception is raised.
"""
import ctypes as C
import os

from . import batch_config as bcfg

_HERE = os.path.dirname(os.path.abspath(__file__))
# FFB_LIB_PATH: load another build of the same sources (kernel A/B experiments, trace builds)
LIB_PATH = os.environ.get("FFB_LIB_PATH") or os.path.join(_HERE, "libffb200.so")


class FFBError(RuntimeError):
    pass


class AttnDesc(C.Structure):
    _fields_ = [
        ("mode", C.c_int), ("dtype", C.c_int), ("num_q_heads", C.c_int), ("num_kv_heads", C.c_int),
        ("head_dim", C.c_int), ("global_num_q_heads", C.c_int),
        ("apply_rotary_embedding", C.c_int), ("rope_theta", C.c_float), ("rope_llama3", C.c_int),
        ("rope_factor", C.c_float), ("rope_low_freq_factor", C.c_float), ("rope_high_freq_factor", C.c_float),
        ("rope_original_max_position_embeddings", C.c_int),
        ("scaling_query", C.c_int), ("scaling_factor", C.c_float), ("qk_prod_scaling", C.c_int),
        ("position_bias", C.c_int),
        ("max_requests", C.c_int), ("max_sequence_length", C.c_int), ("max_spec_tree_tokens", C.c_int),
        ("max_tokens_per_batch", C.c_int),
    ]


class RMDesc(C.Structure):
    _fields_ = [("max_requests_per_batch", C.c_int), ("max_tokens_per_batch", C.c_int),
                ("max_spec_tree_token_num", C.c_int), ("max_sequence_length", C.c_int), ("num_ssms", C.c_int),
                ("num_eos_tokens", C.c_int), ("eos_token_ids", C.c_int * 8)]


class LinearDesc(C.Structure):
    _fields_ = [("dtype", C.c_int), ("quant_type", C.c_int), ("in_dim", C.c_int), ("out_dim", C.c_int),
                ("max_batch", C.c_int), ("activation", C.c_int)]


# name -> (restype, argtypes); every symbol include/ffb200.h declares
_vp, _i, _sz = C.c_void_p, C.c_int, C.c_size_t
SIGNATURES = {
    "ffb_last_error": (C.c_char_p, []),
    "ffb_version": (C.c_char_p, []),
    "ffb_check_device": (_i, []),
    "ffb_reload_env": (None, []),
    "ffb_attn_create": (_i, [C.POINTER(AttnDesc), C.POINTER(_vp)]),
    "ffb_attn_destroy": (_i, [_vp]),
    "ffb_attn_device_bytes": (_sz, [_vp]),
    "ffb_attn_set_profiling": (_i, [_vp, _i]),
    "ffb_attn_last_kernel_ms": (_i, [_vp, C.POINTER(C.c_float)]),
    "ffb_load_batch_config": (_i, [_vp, _i, _vp, _vp, _vp]),
    "ffb_inc_attn_forward": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "ffb_tree_attn_forward": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "ffb_spec_attn_forward": (_i, [_vp, _vp, _vp, _i, _vp, _vp, _vp]),
    "ffb_attn_read_kv": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "ffb_attn_write_kv": (_i, [_vp, _i, _i, _i, _vp, _vp, _vp]),
    "ffb_quant_blob_bytes": (_sz, [_i, _i, _i, _i]),
    "ffb_dequant_int4": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "ffb_dequant_int8": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "ffb_linear_create": (_i, [C.POINTER(LinearDesc), C.POINTER(_vp)]),
    "ffb_linear_destroy": (_i, [_vp]),
    "ffb_linear_forward": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "ffb_linear_forward_static": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "ffb_linear_static_scratch_bytes": (C.c_size_t, []),
    "ffb_comm_get_unique_id": (_i, [_vp]),
    "ffb_comm_init_rank": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "ffb_comm_destroy": (_i, [_vp]),
    "ffb_comm_from_nccl": (_i, [_vp, _i, _i, C.POINTER(_vp)]),
    "ffb_allreduce": (_i, [_vp, _vp, _vp, _sz, _i, _vp]),
    "ffb_comm_has_peer": (_i, [_vp]),
    "ffb_linear_forward_allreduce": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i, _vp]),
    "ffb_residual_rms_norm": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, C.c_float, _i, _vp]),
    "ffb_argmax": (_i, [_vp, _vp, _vp, _vp, _i, _i, _i, _vp]),
    "ffb_arg_topk": (_i, [_vp, _vp, _vp, _i, _i, _i, _i, _vp]),
    "ffb_softmax": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "ffb_bitmask_init": (None, [_vp, _i]),
    "ffb_bitmask_update": (None, [_vp, _i, _i]),
    "ffb_bitmask_append_pending": (None, [_vp, _i]),
    "ffb_bitmask_append": (_i, [_vp, _i, _vp, _i]),
    "ffb_traverse_verify_tree": (_i, [_vp, _i, _vp, _i, _vp, _vp, C.POINTER(_i), _vp, C.POINTER(_i)]),
    "ffb_rm_create": (_i, [C.POINTER(RMDesc), C.POINTER(_vp)]),
    "ffb_rm_destroy": (_i, [_vp]),
    "ffb_rm_push_spec_infer_tree_width": (_i, [_vp, _i]),
    "ffb_rm_register_request": (_i, [_vp, _vp, _i, _i, _i, C.POINTER(_sz)]),
    "ffb_rm_prepare_next_batch": (_i, [_vp, _vp, _vp, _vp]),
    "ffb_rm_prepare_next_batch_init": (_i, [_vp, _vp, _vp, _i, _vp]),
    "ffb_rm_prepare_next_batch_beam": (_i, [_vp, _vp, _vp, _vp]),
    "ffb_rm_prepare_next_batch_verify": (_i, [_vp, _vp, _i, _vp]),
    "ffb_rm_request_status": (_i, [_vp, _sz, C.POINTER(_i)]),
    "ffb_rm_request_tokens": (_i, [_vp, _sz, _vp, _i, C.POINTER(_i)]),
    "ffb_rm_request_steps": (_i, [_vp, _sz, C.POINTER(_i), C.POINTER(_i)]),
    "ffb_rm_num_pending": (_i, [_vp]),
    "ffb_rm_num_processed_requests": (_sz, [_vp]),
    "ffb_fetch_inference_result": (_i, [_vp, _i, _vp, _vp]),
    "ffb_load_from_file": (_i, [C.c_char_p, _vp, _sz, _i]),
    "ffb_load_attention_weights_to_dense": (_i, [C.c_char_p, C.c_char_p, _i, _i, _i, _i, _i, _i, _i, _i, _vp]),
    "ffb_quantize_weights": (_i, [_vp, _i, _i, _i, _i, _vp]),
    "ffb_save_quantized_file": (_i, [C.c_char_p, _vp, _i, _i, _i, _i]),
    "ffb_load_from_quantized_file": (_i, [C.c_char_p, _i, _i, _sz, _vp]),
    "ffb_fetch_beam_inference_result": (_i, [_vp, _vp, _i, _i, _vp, _vp]),
}

_lib = None


def load():
    """Load the CUDA extension; raises FFBError when it has not been built (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FFBError("%s is missing: run `python -m flexflow_b200.build` (nvcc, sm_100a). "
                       "There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def reload_env():
    """Have the library read its FFB_* override variables again (they are cached at first use)."""
    load().ffb_reload_env()


def check(rc, what=""):
    if rc != 0:
        msg = load().ffb_last_error()
        raise FFBError("%s failed with code %d: %s" % (what or "ffb call", rc, (msg or b"").decode(errors="replace")))
