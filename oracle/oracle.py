"""Python face of the CPU oracle (oracle/*.cpp).  TEST INFRASTRUCTURE ONLY: imported by tests/,
__graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs, never by flexflow_b200/."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liborc.so")
_lib = None

DT_CODE = {"float": 0, "half": 1, "bf16": 2}


def build(force=False):
    srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith((".cpp", ".h"))]
    srcs.append(os.path.join(_HERE, "..", "include", "ffb200_batch_config.h"))
    if force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs):
        subprocess.run(["make", "-C", _HERE, "-s", "-B"], check=True)
    return _SO


_REF_SO = os.path.join(_HERE, "_ref", "libref_decompress.so")
_ref = None


def build_ref():
    """Compile the reference's own decompression kernels (oracle/Makefile target `ref`) where /root/reference is
    mounted; elsewhere (the GPU box) the prebuilt oracle/_ref/*.so that travelled with the snapshot is used."""
    if os.path.isdir("/root/reference/src/ops/kernels"):
        subprocess.run(["make", "-C", _HERE, "-s", "ref"], check=True)
    return _REF_SO if os.path.exists(_REF_SO) else None


def ref_lib():
    """The reference's decompress kernels compiled for sm_100a (GPU only), or None when the .so was not built."""
    global _ref
    if _ref is None and os.path.exists(_REF_SO):
        _ref = C.CDLL(_REF_SO)
        for n in ("ref_decompress_int4_half", "ref_decompress_int8_half"):
            getattr(_ref, n).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    return _ref


_REF_ATTN_SO = os.path.join(_HERE, "_ref", "libref_attn.so")
_ref_attn = None


def ref_attn_lib():
    """The reference's decode attention kernel (inc_multihead_self_attention.cu:372-623) compiled for sm_100a, or None."""
    global _ref_attn
    if _ref_attn is None and os.path.exists(_REF_ATTN_SO):
        _ref_attn = C.CDLL(_REF_ATTN_SO)
        _ref_attn.ref_inc_generation_half.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                                        C.c_int, C.c_int, C.c_void_p]
        _ref_attn.ref_spec_generation_half.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_int, C.c_int, C.c_int] + \
            [C.c_void_p] * 4 + [C.c_int, C.c_int, C.c_void_p]
        _ref_attn.ref_tree_verify_half.argtypes = [C.c_void_p] * 4 + [C.c_float, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                                                     C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int, C.c_int,
                                                                     C.c_void_p]
        _ref_attn.ref_commit_tokens_half.argtypes = [C.c_void_p] * 4 + [C.c_int] * 5 + [C.c_void_p]
        _ref_attn.ref_update_tree_branch_kv_cache_half.argtypes = [C.c_void_p] * 5 + [C.c_int] * 4 + [C.c_void_p]
        _ref_attn.ref_apply_rotary_embedding_half.argtypes = [C.c_void_p] * 3 + [C.c_float, C.c_int, C.c_float, C.c_float,
                                                                                 C.c_float, C.c_int, C.c_int, C.c_int,
                                                                                 C.c_int, C.c_void_p]
    return _ref_attn


_REF_NORM_SO = os.path.join(_HERE, "_ref", "libref_norm.so")
_ref_norm = None


def ref_norm_lib():
    """The reference's (Residual)RMSNorm forward kernels and ArgMax's cub call compiled for sm_100a, or None."""
    global _ref_norm
    if _ref_norm is None and os.path.exists(_REF_NORM_SO):
        _ref_norm = C.CDLL(_REF_NORM_SO)
        _ref_norm.ref_residual_rms_norm_half.argtypes = [C.c_void_p] * 7 + [C.c_int, C.c_int, C.c_float, C.c_void_p]
        _ref_norm.ref_rms_norm_half.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_float, C.c_void_p]
        _ref_norm.ref_arg_topk_half.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_void_p]
        _ref_norm.ref_argmax_half.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p,
                                              C.c_size_t, C.POINTER(C.c_size_t), C.c_void_p, C.c_void_p]
    return _ref_norm


_REF_LINEAR_SO = os.path.join(_HERE, "_ref", "libref_linear.so")
_ref_linear = None


def ref_linear_lib():
    """The reference Linear's cublasGemmEx call (fp16 accumulation) restated, or None when the .so was not built."""
    global _ref_linear
    if _ref_linear is None and os.path.exists(_REF_LINEAR_SO):
        try:
            _ref_linear = C.CDLL(_REF_LINEAR_SO)
        except OSError:  # no cuBLAS on this machine
            return None
        _ref_linear.ref_linear_forward_half.argtypes = [C.c_void_p] * 3 + [C.c_int] * 3 + [C.c_void_p]
    return _ref_linear


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.orc_attn_create.restype = C.c_void_p
        _lib.orc_attn_create.argtypes = [C.c_int] * 9 + [C.c_float, C.c_int, C.c_float, C.c_float, C.c_float, C.c_int,
                                                          C.c_int, C.c_float, C.c_int, C.c_int, C.c_int]
        _lib.orc_attn_destroy.argtypes = [C.c_void_p]
        for n in ("orc_attn_kcache", "orc_attn_vcache"):
            getattr(_lib, n).restype = C.POINTER(C.c_float)
            getattr(_lib, n).argtypes = [C.c_void_p]
        _lib.orc_attn_cache_len.argtypes = [C.c_void_p]
        _lib.orc_attn_set_quirks.argtypes = [C.c_void_p, C.c_int]
        for n in ("orc_inc_attn_forward", "orc_tree_attn_forward", "orc_spec_attn_forward"):
            getattr(_lib, n).argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
        _lib.orc_dequant_int4.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        _lib.orc_dequant_int8.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int]
        _lib.orc_quantize.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib.orc_linear_forward.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int]
        _lib.orc_set_threads.argtypes = [C.c_int]
        _lib.orc_round_to_dt.argtypes = [C.c_void_p, C.c_size_t, C.c_int]
        _lib.orc_float_to_half_bits.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib.orc_half_bits_to_float.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t]
        _lib.orc_init_bitmask.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_update_bitmask.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int]
        _lib.orc_append_pending_request.argtypes = [C.c_void_p, C.c_int]
        _lib.orc_append_bitmask.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
        _lib.orc_residual_rms_norm.argtypes = [C.c_void_p] * 5 + [C.c_int, C.c_int, C.c_float, C.c_int]
        _lib.orc_argmax.argtypes = [C.c_void_p] * 4 + [C.c_int, C.c_int]
        _lib.orc_arg_topk.argtypes = [C.c_void_p] * 3 + [C.c_int, C.c_int, C.c_int]
        _lib.orc_traverse_verify_tree.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_int] + [C.c_void_p] * 9
    return _lib


def set_threads(n):
    lib().orc_set_threads(int(n))


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def round_to_dt(x, dt):
    x = np.ascontiguousarray(x, dtype=np.float32).copy()
    lib().orc_round_to_dt(_p(x), x.size, DT_CODE[dt])
    return x


class OracleAttn:
    """State of one attention operator in the oracle (role of …Meta): reference-layout caches."""

    def __init__(self, mode, dt, num_q_heads, head_dim, max_requests, max_seq, max_spec, max_tokens,
                 apply_rope=True, theta=10000.0, llama3=False, factor=8.0, low_ff=1.0, high_ff=4.0,
                 orig_max_pos=8192, scaling_query=False, scaling_factor=1.0, qk_prod_scaling=True,
                 position_bias=False, global_nq=None):
        self.nq, self.dh, self.H = num_q_heads, head_dim, num_q_heads * head_dim
        self.max_requests, self.dt = max_requests, dt
        self._h = lib().orc_attn_create(mode, DT_CODE[dt], num_q_heads, head_dim, max_requests, max_seq, max_spec,
                                        max_tokens, int(apply_rope), theta, int(llama3), factor, low_ff, high_ff,
                                        orig_max_pos, int(scaling_query), scaling_factor, int(qk_prod_scaling),
                                        int(position_bias), global_nq or num_q_heads)
        self.cache_len = lib().orc_attn_cache_len(self._h)
        shape = (max_requests, self.cache_len, self.H)
        n = int(np.prod(shape))
        self.kcache = np.ctypeslib.as_array(lib().orc_attn_kcache(self._h), shape=(n,)).reshape(shape)
        self.vcache = np.ctypeslib.as_array(lib().orc_attn_vcache(self._h), shape=(n,)).reshape(shape)

    def set_quirks(self, on):
        lib().orc_attn_set_quirks(self._h, int(on))

    def _fwd(self, name, bc, shard_id, qkv):
        T = bc.num_tokens
        qkv = np.ascontiguousarray(qkv, dtype=np.float32)
        assert qkv.shape == (T, 3, self.H), qkv.shape
        out = np.zeros((T, self.H), dtype=np.float32)
        rc = getattr(lib(), name)(self._h, C.addressof(bc), shard_id, _p(qkv), _p(out))
        if rc != 0:
            raise RuntimeError("%s violated a reference assert (code %d)" % (name, rc))
        return out

    def inc_forward(self, bc, qkv, shard_id=0):
        return self._fwd("orc_inc_attn_forward", bc, shard_id, qkv)

    def tree_forward(self, bc, qkv, shard_id=0):
        return self._fwd("orc_tree_attn_forward", bc, shard_id, qkv)

    def spec_forward(self, bc, qkv, shard_id=0):
        return self._fwd("orc_spec_attn_forward", bc, shard_id, qkv)

    def __del__(self):
        try:
            lib().orc_attn_destroy(self._h)
        except Exception:
            pass


def blob_bytes(bits, in_dim, out_dim):
    n = in_dim * out_dim
    return (n // 2 if bits == 4 else n) + 2 * (n // 32) * 2


def quantize(w, bits, dt="half"):
    """[out][in] float32 -> FlexFlow blob [values | offsets | scales] (uint8 array)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    out_dim, in_dim = w.shape
    blob = np.zeros(blob_bytes(bits, in_dim, out_dim), dtype=np.uint8)
    lib().orc_quantize(_p(w), _p(blob), in_dim, out_dim, bits, DT_CODE[dt])
    return blob


def dequant(blob, bits, in_dim, out_dim, dt="half"):
    """Returns the dequantised [out][in] matrix as uint16 bit patterns of DT."""
    out = np.zeros((out_dim, in_dim), dtype=np.uint16)
    fn = lib().orc_dequant_int4 if bits == 4 else lib().orc_dequant_int8
    fn(_p(np.ascontiguousarray(blob)), _p(out), in_dim, out_dim, DT_CODE[dt])
    return out


def linear(x, w, bias=None, dt="half"):
    x = np.ascontiguousarray(x, dtype=np.float32)
    w = np.ascontiguousarray(w, dtype=np.float32)
    out = np.zeros((x.shape[0], w.shape[0]), dtype=np.float32)
    b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
    lib().orc_linear_forward(_p(x), _p(w), _p(b) if b is not None else None, _p(out), x.shape[0], x.shape[1],
                             w.shape[0], DT_CODE[dt])
    return out


def to_bits(x, dt="half"):
    """float array -> uint16 bit patterns of DT (round to nearest even)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    if dt == "half":
        out = np.zeros(x.shape, dtype=np.uint16)
        lib().orc_float_to_half_bits(_p(x), _p(out), x.size)
        return out
    r = round_to_dt(x, "bf16")
    return (r.view(np.uint32) >> 16).astype(np.uint16)


def from_bits(b, dt="half"):
    b = np.ascontiguousarray(b, dtype=np.uint16)
    if dt == "half":
        out = np.zeros(b.shape, dtype=np.float32)
        lib().orc_half_bits_to_float(_p(b), _p(out), b.size)
        return out
    return (b.astype(np.uint32) << 16).view(np.float32)


def residual_rms_norm(x1_bits, x2_bits, weight_bits, eps, dt="half"):
    """16-bit patterns in and out.  Returns (residual_output_bits or None, output_bits); x2_bits None = RMSNorm."""
    x1 = np.ascontiguousarray(x1_bits, dtype=np.uint16)
    rows, n = x1.shape
    w = np.ascontiguousarray(weight_bits, dtype=np.uint16)
    out = np.zeros_like(x1)
    if x2_bits is None:
        lib().orc_residual_rms_norm(_p(x1), None, _p(w), None, _p(out), rows, n, eps, DT_CODE[dt])
        return None, out
    x2 = np.ascontiguousarray(x2_bits, dtype=np.uint16)
    res = np.zeros_like(x1)
    lib().orc_residual_rms_norm(_p(x1), _p(x2), _p(w), _p(res), _p(out), rows, n, eps, DT_CODE[dt])
    return res, out


def argmax(x, beam_search=False):
    """x: [rows][length] float32 (values already in DT).  Returns (indices, prob or None, parent or None)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows, length = x.shape
    idx = np.zeros(rows, dtype=np.int32)
    prob = np.zeros(rows, dtype=np.float32) if beam_search else None
    parent = np.full(rows, -1, dtype=np.int32) if beam_search else None
    lib().orc_argmax(_p(x), _p(idx), _p(prob) if beam_search else None, _p(parent) if beam_search else None, rows,
                     length)
    return idx, prob, parent


def arg_topk(x, k):
    """x: [rows][length] float32.  Returns (indices [rows][k], values [rows][k] with the reference's unwritten last
    column set to NaN)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    rows, length = x.shape
    idx = np.zeros((rows, k), dtype=np.int32)
    val = np.full((rows, k), np.nan, dtype=np.float32)
    lib().orc_arg_topk(_p(x), _p(val), _p(idx), rows, length, k)
    return idx, val


def traverse_verify_tree(in_tree, out_tree, committed):
    """in_tree/out_tree: lists of (token, depth); committed: list of (depth, index_in_batch).
    Returns (verified [(token, depth)], new_committed [(depth, index)])."""
    n_in, n_out = len(in_tree), len(out_tree)
    arr = lambda xs: np.ascontiguousarray(np.array(xs, dtype=np.int32))
    it, idp = arr([t for t, _ in in_tree]), arr([d for _, d in in_tree])
    ot, odp = arr([t for t, _ in out_tree]), arr([d for _, d in out_tree])
    cd, ci = arr([d for d, _ in committed]), arr([i for _, i in committed])
    vt, vd = np.zeros(n_in + 1, np.int32), np.zeros(n_in + 1, np.int32)
    nd, ni = np.zeros(n_in + 1, np.int32), np.zeros(n_in + 1, np.int32)
    nc = C.c_int(0)
    nv = lib().orc_traverse_verify_tree(n_in, _p(it), _p(idp), n_out, _p(ot), _p(odp), _p(cd), _p(ci), _p(vt), _p(vd),
                                        _p(nd), _p(ni), C.byref(nc))
    if nv < 0:
        raise RuntimeError("traverse_verify_tree violated a reference assert (%d)" % nv)
    return [(int(vt[i]), int(vd[i])) for i in range(nv)], [(int(nd[i]), int(ni[i])) for i in range(nc.value)]
