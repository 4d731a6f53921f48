"""CPU: the reference's weight file formats and quantisation rule behind the C ABI (flexflow_b200/csrc/weights_io.cu,
SURVEY §8f N4), pinned against reference code:
  * the quantiser against tests/golden/ref_compress.npz, produced by compress() imported from the reference's
    inference/utils/compress_llama_weights.py (generator committed: tests/golden/make_compress_golden.py);
  * the file readers against the reference's own load_attention_weights_to_dense_v2 / load_from_file /
    load_from_quantized_file compiled from src/runtime/file_loader.cc into oracle/_ref/libref_loader.so."""
import ctypes as C
import os

import numpy as np
import pytest

from flexflow_b200 import batch_config as B
from flexflow_b200 import file_loader as FL
from flexflow_b200 import lib as L
from oracle import oracle as O

HERE = os.path.dirname(os.path.abspath(__file__))
_REF_SO = os.path.join(HERE, "..", "oracle", "_ref", "libref_loader.so")


def _ref():
    if not os.path.exists(_REF_SO):
        pytest.skip("oracle/_ref/libref_loader.so not built (make -C oracle ref)")
    R = C.CDLL(_REF_SO)
    vp, i, sz = C.c_void_p, C.c_int, C.c_size_t
    R.ref_load_attention_weights_to_dense_v2.argtypes = [vp, i, i, i, sz, sz, C.c_char_p, C.c_char_p, sz, i, i]
    R.ref_load_from_file.argtypes = [vp, i, sz, C.c_char_p]
    R.ref_load_from_quantized_file.argtypes = [vp, sz, C.c_char_p, i, i]
    return R


def _sections(blob, bits, in_dim, out_dim):
    n = in_dim * out_dim
    nv = n // 2 if bits == 4 else n
    nm = n // 32 * 2
    vals, off, sc = blob[:nv], blob[nv:nv + nm].view(np.uint16), blob[nv + nm:].view(np.uint16)
    if bits == 4:
        codes = np.empty(n, np.uint8)
        codes[0::2], codes[1::2] = vals >> 4, vals & 0xF  # element 2b in the high nibble (decompress_kernels.cu:43-50)
    else:
        codes = vals
    return codes.reshape(out_dim, in_dim), off.reshape(out_dim // 32, in_dim), sc.reshape(out_dim // 32, in_dim)


@pytest.mark.parametrize("bits", [8, 4])
def test_quantiser_equals_the_reference_compress(bits):
    fx = np.load(os.path.join(HERE, "golden", "ref_compress.npz"))
    for name in "abc":
        w = fx["w_" + name]
        out_dim, in_dim = w.shape
        codes, off, sc = _sections(FL.compress(w, bits), bits, in_dim, out_dim)
        assert np.array_equal(codes, fx["q%d_%s" % (bits, name)]), name
        assert np.array_equal(off, fx["mn%d_%s" % (bits, name)].astype(np.float16).view(np.uint16))
        assert np.array_equal(sc, fx["sc%d_%s" % (bits, name)].astype(np.float16).view(np.uint16))


@pytest.mark.parametrize("bits", [8, 4])
def test_quantiser_equals_the_oracle_and_survives_constant_groups(bits):
    rng = np.random.default_rng(3)
    w = (rng.standard_normal((128, 96)) * 0.05).astype(np.float32)
    assert np.array_equal(FL.compress(w, bits), O.quantize(w, bits))
    w[32:64, 5] = 0.125  # a constant (32-row group, column): q = 0, scale = inf, dequantises to the constant
    blob = FL.compress(w, bits)
    codes, off, sc = _sections(blob, bits, 96, 128)
    assert (codes[32:64, 5] == 0).all() and sc[1, 5] == 0x7C00 and off[1, 5] == np.float16(0.125).view(np.uint16)
    dq = O.from_bits(O.dequant(blob, bits, 96, 128))
    assert (dq[32:64, 5] == 0.125).all()
    step = (w.reshape(4, 32, 96).max(1) - w.reshape(4, 32, 96).min(1)) / (2 ** bits - 1)
    assert (np.abs(dq - w).reshape(4, 32, 96) <= step[:, None, :] * 0.51 + 2e-3 * np.abs(w).max()).all()


@pytest.mark.parametrize("bits", [8, 4])
def test_quantised_files_round_trip_and_equal_the_reference_reader(bits, tmp_path):
    R = _ref()
    rng = np.random.default_rng(bits)
    in_dim, out_dim = 80, 160
    w = rng.standard_normal((out_dim, in_dim)).astype(np.float32)
    blob = FL.compress(w, bits)
    path = str(tmp_path / "layers.0.mlp.gate_proj.weight")
    FL.save_quantized(path, blob, in_dim, out_dim, bits)
    # the value file holds one byte per element, also for 4-bit codes (file_loader.cc:596-612)
    assert os.path.getsize(path) == in_dim * out_dim
    assert os.path.getsize(path + "_offset") == os.path.getsize(path + "_scale") == in_dim * out_dim // 32 * 2
    ld = FL.FileDataLoader(str(tmp_path), 1, 1, in_dim, 1)
    qt = B.DT_INT4 if bits == 4 else B.DT_INT8
    back = ld.load_from_quantized_file("layers.0.mlp.gate_proj.weight", in_dim, out_dim, qt)
    assert np.array_equal(back, blob)
    ref = np.zeros_like(blob)
    R.ref_load_from_quantized_file(ref.ctypes.data_as(C.c_void_p), blob.size, path.encode(), qt, 0)
    assert np.array_equal(ref, blob)


@pytest.mark.parametrize("nh,nkv,tp", [(8, 8, 1), (8, 8, 4), (8, 2, 2), (8, 2, 4), (16, 4, 8), (4, 1, 2)])
@pytest.mark.parametrize("dtype", ["half", "float"])
def test_dense_attention_loader_equals_the_reference_loader(nh, nkv, tp, dtype, tmp_path):
    R = _ref()
    import torch
    from flexflow_b200 import sharding
    rng = np.random.default_rng(nh * 100 + nkv * 10 + tp)
    hidden, dh = 48, 8
    npdt = np.uint16 if dtype == "half" else np.float32
    es = 2 if dtype == "half" else 4

    def rnd(shape):
        a = rng.integers(0, 60000, size=shape).astype(npdt) if dtype == "half" else rng.standard_normal(shape).astype(np.float32)
        return a
    wq, wk, wv, wo = rnd((nh * dh, hidden)), rnd((nkv * dh, hidden)), rnd((nkv * dh, hidden)), rnd((hidden, nh * dh))
    name = "layers.3.self_attn"
    for suffix, a in ((".q_proj.weight", wq), (".k_proj.weight", wk), (".v_proj.weight", wv), (".o_proj.weight", wo)):
        a.tofile(str(tmp_path / (name + suffix)))
    ld = FL.FileDataLoader(str(tmp_path), nh, nkv, hidden, dh, tp, use_full_precision=(dtype == "float"))
    full = ld.load_attention_weights_to_dense(name, False, tp_rank=-1)
    ref = np.zeros_like(full)
    R.ref_load_attention_weights_to_dense_v2(ref.ctypes.data_as(C.c_void_p), es, nh, nkv, hidden, dh, name.encode(),
                                             str(tmp_path).encode(), full.size, tp, 0)
    assert np.array_equal(full, ref)
    full_o = ld.load_attention_weights_to_dense(name, True, tp_rank=-1)
    ref_o = np.zeros_like(full_o)
    R.ref_load_attention_weights_to_dense_v2(ref_o.ctypes.data_as(C.c_void_p), es, nh, nkv, hidden, dh, name.encode(),
                                             str(tmp_path).encode(), full_o.size, tp, 1)
    assert np.array_equal(full_o, ref_o) and np.array_equal(full_o, wo)
    H_l = nh // tp * dh
    as_t = (lambda a: torch.from_numpy(a.view(np.int16))) if dtype == "half" else torch.from_numpy
    for r in range(tp):
        shard = ld.load_attention_weights_to_dense(name, False, tp_rank=r)
        assert np.array_equal(shard, full[r * 3 * H_l:(r + 1) * 3 * H_l])       # Legion hands shard r this slice
        want = sharding.shard_qkv_proj(as_t(wq), as_t(wk), as_t(wv), nh, nkv, dh, tp, r).numpy()
        assert np.array_equal(shard.view(want.dtype), want)
        o_shard = ld.load_attention_weights_to_dense(name, True, tp_rank=r)
        assert np.array_equal(o_shard, wo[:, r * H_l:(r + 1) * H_l])
    # plain tensors (norm weights, lm_head, ...)
    norm = rnd((hidden,))
    norm.tofile(str(tmp_path / "layers.3.input_layernorm.weight"))
    got = ld.load_from_file("layers.3.input_layernorm.weight", hidden)
    ref_n = np.zeros_like(got)
    R.ref_load_from_file(ref_n.ctypes.data_as(C.c_void_p), es, hidden, str(tmp_path / "layers.3.input_layernorm.weight").encode())
    assert np.array_equal(got, norm) and np.array_equal(ref_n, norm)


def test_missing_and_short_files_are_reported(tmp_path):
    ld = FL.FileDataLoader(str(tmp_path), 4, 4, 16, 4)
    with pytest.raises(L.FFBError):
        ld.load_from_file("nope.weight", 16)
    np.zeros(10, np.uint16).tofile(str(tmp_path / "short.weight"))
    with pytest.raises(L.FFBError):
        ld.load_from_file("short.weight", 16)
    with pytest.raises(L.FFBError):
        ld.load_attention_weights_to_dense("layers.0.self_attn", False)
    with pytest.raises(L.FFBError):
        FL.compress(np.zeros((40, 8), np.float32), 8)  # out_dim not a multiple of the group size
