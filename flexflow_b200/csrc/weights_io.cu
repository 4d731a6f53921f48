// weights_io.cu — the reference's weight file formats and its quantisation rule behind the C ABI (host code):
//   * HF-export binaries: one flat file of DT elements per parameter; the q / k / v files of a layer are fused into the
//     column-parallel qkv_proj weight [Q_r | K_r | V_r] per TP shard with the kv heads replicated to q-head
//     granularity, o_proj is taken as stored (FileDataLoader, src/runtime/file_loader.cc:215-377);
//   * quantised tensors: `<name>` holds one byte per element (also for 4-bit values), `<name>_offset` and
//     `<name>_scale` one DT per (32-row group, column); in memory the three become the blob [values | offsets | scales]
//     with two 4-bit values per byte, the FIRST in the high nibble (load_from_quantized_file, file_loader.cc:552-669);
//   * the quantiser of inference/utils/compress_llama_weights.py:15-58 (asymmetric, groups of 32 along the output
//     dimension): scale = B / (max - min), q = round((w - min) * scale) clamped to [0, B], fp32 arithmetic - the shipped
//     script stops before its save step; this one writes the three files the loader expects.
// The blob is what ffb_dequant_int4 / _int8 and the quantised Linear paths consume.
#include <algorithm>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

#include "common.cuh"

namespace {

int elt_size(int dtype) { return dtype == FFB_DT_FLOAT ? 4 : ((dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) ? 2 : 0); }

std::string join2(char const *folder, std::string const &name) { // join_path, src/runtime/inference_manager.cc:824-841
  std::string j = folder ? folder : "";
  if (!name.empty() && name[0] == '/') {
    return name;
  }
  if (!j.empty() && j.back() != '/') {
    j += '/';
  }
  return j + name;
}

// reads exactly `bytes` bytes; a short file is an error (the reference asserts)
int read_exact(std::string const &path, void *dst, size_t bytes) {
  FILE *f = fopen(path.c_str(), "rb");
  if (!f) {
    ffb::set_error("could not open file: %s", path.c_str());
    return FFB_ERR_INVALID;
  }
  size_t const got = fread(dst, 1, bytes, f);
  fclose(f);
  if (got != bytes) {
    ffb::set_error("%s holds %zu bytes, %zu expected", path.c_str(), got, bytes);
    return FFB_ERR_INVALID;
  }
  return FFB_OK;
}

int write_all(std::string const &path, void const *src, size_t bytes) {
  FILE *f = fopen(path.c_str(), "wb");
  if (!f) {
    ffb::set_error("could not create file: %s", path.c_str());
    return FFB_ERR_INVALID;
  }
  size_t const put = fwrite(src, 1, bytes, f);
  int const rc = fclose(f);
  if (put != bytes || rc != 0) {
    ffb::set_error("short write to %s", path.c_str());
    return FFB_ERR_INVALID;
  }
  return FFB_OK;
}

// round-to-nearest-even conversions of the group meta data (the same bit patterns cuda_fp16 / cuda_bf16 produce)
uint16_t to_meta_bits(float v, int dtype) {
  if (dtype == FFB_DT_BF16) {
    __nv_bfloat16 b = __float2bfloat16_rn(v);
    uint16_t u;
    memcpy(&u, &b, 2);
    return u;
  }
  __half h = __float2half_rn(v);
  uint16_t u;
  memcpy(&u, &h, 2);
  return u;
}

struct BlobShape {
  size_t values, meta; // bytes of the value section, bytes of EACH of the two meta sections
};
bool blob_shape(int quant_type, int dtype, size_t in_dim, size_t out_dim, BlobShape *s) {
  if ((quant_type != FFB_DT_INT4 && quant_type != FFB_DT_INT8) || elt_size(dtype) != 2 || out_dim % 32 != 0 ||
      in_dim == 0 || out_dim == 0 || (quant_type == FFB_DT_INT4 && (in_dim * out_dim) % 2 != 0)) {
    return false;
  }
  size_t const n = in_dim * out_dim;
  s->values = quant_type == FFB_DT_INT4 ? n / 2 : n;
  s->meta = n / 32 * 2;
  return true;
}

} // namespace

extern "C" {

int ffb_load_from_file(char const *path, void *dst_host, size_t count, int dtype) {
  FFB_REQUIRE(path && dst_host && elt_size(dtype) > 0, "bad argument");
  return read_exact(path, dst_host, count * (size_t)elt_size(dtype));
}

int ffb_load_attention_weights_to_dense(char const *weights_folder, char const *layer_name, int dtype, int num_heads,
                                        int num_kv_heads, int hidden_dim, int head_dim, int tp_degree, int tp_rank,
                                        int load_o_proj, void *dst_host) {
  FFB_REQUIRE(weights_folder && layer_name && dst_host, "null argument");
  size_t const es = (size_t)elt_size(dtype);
  FFB_REQUIRE(es > 0 && num_heads > 0 && num_kv_heads > 0 && num_heads % num_kv_heads == 0 && hidden_dim > 0 && head_dim > 0 &&
                  tp_degree >= 1 && num_heads % tp_degree == 0 && tp_rank >= -1 && tp_rank < tp_degree,
              "bad shape / sharding arguments");
  size_t const head_elems = (size_t)hidden_dim * head_dim;  // one head's [head_dim][hidden] rows (or o_proj columns)
  size_t const head_bytes = head_elems * es;
  int const heads_per_shard = num_heads / tp_degree, group = num_heads / num_kv_heads;
  int const r0 = tp_rank < 0 ? 0 : tp_rank, r1 = tp_rank < 0 ? tp_degree : tp_rank + 1;
  char *dst = (char *)dst_host;
  std::string const base = layer_name;
  if (load_o_proj) {
    // stored [hidden][num_heads * head_dim]; the row-parallel shard r owns the columns of its heads
    size_t const in_full = (size_t)num_heads * head_dim, in_shard = (size_t)heads_per_shard * head_dim;
    std::vector<char> file((size_t)hidden_dim * in_full * es);
    int rc = read_exact(join2(weights_folder, base + ".o_proj.weight"), file.data(), file.size());
    if (rc != FFB_OK) {
      return rc;
    }
    if (tp_rank < 0) {
      memcpy(dst, file.data(), file.size());
      return FFB_OK;
    }
    for (int o = 0; o < hidden_dim; o++) {
      memcpy(dst + (size_t)o * in_shard * es, file.data() + ((size_t)o * in_full + (size_t)tp_rank * in_shard) * es, in_shard * es);
    }
    return FFB_OK;
  }
  // q file: num_heads heads; k, v files: num_kv_heads heads.  Shard r of the fused weight is
  // [Q heads of r | K head of each of those q heads | V likewise], every head a block of head_elems.
  char const *const suffix[3] = {".q_proj.weight", ".k_proj.weight", ".v_proj.weight"};
  size_t const shard_bytes = (size_t)3 * heads_per_shard * head_bytes;
  for (int part = 0; part < 3; part++) {
    int const file_heads = part == 0 ? num_heads : num_kv_heads;
    std::vector<char> file((size_t)file_heads * head_bytes);
    int rc = read_exact(join2(weights_folder, base + suffix[part]), file.data(), file.size());
    if (rc != FFB_OK) {
      return rc;
    }
    for (int r = r0; r < r1; r++) {
      char *out = dst + (size_t)(r - r0) * shard_bytes + (size_t)part * heads_per_shard * head_bytes;
      for (int hl = 0; hl < heads_per_shard; hl++) {
        int const head = r * heads_per_shard + hl;
        int const src_head = part == 0 ? head : head / group;
        memcpy(out + (size_t)hl * head_bytes, file.data() + (size_t)src_head * head_bytes, head_bytes);
      }
    }
  }
  return FFB_OK;
}

int ffb_quantize_weights(float const *w_host, int in_dim, int out_dim, int quant_type, int dtype, void *blob_host) {
  FFB_REQUIRE(w_host && blob_host, "null argument");
  BlobShape bs;
  FFB_REQUIRE(blob_shape(quant_type, dtype, (size_t)in_dim, (size_t)out_dim, &bs),
              "quantisation needs int4 / int8, 16-bit group meta and out_dim %% 32 == 0");
  uint8_t *vals = (uint8_t *)blob_host;
  uint16_t *offs = (uint16_t *)(vals + bs.values), *scales = (uint16_t *)(vals + bs.values + bs.meta);
  float const B = quant_type == FFB_DT_INT4 ? 15.f : 255.f;
  if (quant_type == FFB_DT_INT4) {
    memset(vals, 0, bs.values);
  }
  std::vector<float> mn((size_t)in_dim), mx((size_t)in_dim), sc((size_t)in_dim);
  for (int g = 0; g < out_dim / 32; g++) {
    float const *rows = w_host + (size_t)g * 32 * in_dim;
    std::copy(rows, rows + in_dim, mn.begin());
    std::copy(rows, rows + in_dim, mx.begin());
    for (int r = 1; r < 32; r++) {
      float const *row = rows + (size_t)r * in_dim;
      for (int i = 0; i < in_dim; i++) {
        mn[i] = std::min(mn[i], row[i]);
        mx[i] = std::max(mx[i], row[i]);
      }
    }
    for (int i = 0; i < in_dim; i++) {
      sc[i] = B / (mx[i] - mn[i]); // inf for a constant group: q = 0 and 0 / inf + min reproduces the value
      offs[(size_t)g * in_dim + i] = to_meta_bits(mn[i], dtype);
      scales[(size_t)g * in_dim + i] = to_meta_bits(sc[i], dtype);
    }
    for (int r = 0; r < 32; r++) {
      float const *row = rows + (size_t)r * in_dim;
      size_t const idx0 = (size_t)(g * 32 + r) * in_dim;
      for (int i = 0; i < in_dim; i++) {
        float q = 0.f;
        if (mx[i] != mn[i]) {
          q = std::nearbyint(std::min(std::max((row[i] - mn[i]) * sc[i], 0.f), B)); // clamp, then round half to even
        }
        uint8_t const v = (uint8_t)q;
        size_t const idx = idx0 + i;
        if (quant_type == FFB_DT_INT4) {
          vals[idx >> 1] |= (idx & 1) ? v : (uint8_t)(v << 4);
        } else {
          vals[idx] = v;
        }
      }
    }
  }
  return FFB_OK;
}

int ffb_save_quantized_file(char const *path, void const *blob_host, int in_dim, int out_dim, int quant_type, int dtype) {
  FFB_REQUIRE(path && blob_host, "null argument");
  BlobShape bs;
  FFB_REQUIRE(blob_shape(quant_type, dtype, (size_t)in_dim, (size_t)out_dim, &bs), "bad quantised tensor shape");
  uint8_t const *vals = (uint8_t const *)blob_host;
  std::string const p = path;
  int rc;
  if (quant_type == FFB_DT_INT4) { // the value file holds ONE BYTE per element
    std::vector<uint8_t> wide(bs.values * 2);
    for (size_t b = 0; b < bs.values; b++) {
      wide[2 * b] = vals[b] >> 4;
      wide[2 * b + 1] = vals[b] & 0xF;
    }
    rc = write_all(p, wide.data(), wide.size());
  } else {
    rc = write_all(p, vals, bs.values);
  }
  if (rc == FFB_OK) {
    rc = write_all(p + "_offset", vals + bs.values, bs.meta);
  }
  if (rc == FFB_OK) {
    rc = write_all(p + "_scale", vals + bs.values + bs.meta, bs.meta);
  }
  return rc;
}

int ffb_load_from_quantized_file(char const *path, int quant_type, int dtype, size_t blob_bytes, void *blob_host) {
  FFB_REQUIRE(path && blob_host, "null argument");
  FFB_REQUIRE((quant_type == FFB_DT_INT4 || quant_type == FFB_DT_INT8) && elt_size(dtype) == 2,
              "int4 / int8 values with 16-bit group meta only");
  // section sizes from the blob size, as the reference derives them (file_loader.cc:566-580): 16 (int4) or 32 (int8)
  // value bytes per 2 + 2 meta bytes
  size_t const unit = quant_type == FFB_DT_INT4 ? 20 : 36;
  FFB_REQUIRE(blob_bytes > 0 && blob_bytes % unit == 0, "blob of %zu bytes is not a whole number of 32-element groups", blob_bytes);
  size_t const meta = blob_bytes / unit * 2, values = blob_bytes - 2 * meta;
  uint8_t *vals = (uint8_t *)blob_host;
  std::string const p = path;
  int rc;
  if (quant_type == FFB_DT_INT4) {
    std::vector<uint8_t> wide(values * 2);
    rc = read_exact(p, wide.data(), wide.size());
    if (rc != FFB_OK) {
      return rc;
    }
    for (size_t b = 0; b < values; b++) { // first element in the high nibble
      vals[b] = (uint8_t)((wide[2 * b + 1] & 0xF) | (wide[2 * b] << 4));
    }
  } else {
    rc = read_exact(p, vals, values);
    if (rc != FFB_OK) {
      return rc;
    }
  }
  rc = read_exact(p + "_offset", vals + values, meta);
  if (rc != FFB_OK) {
    return rc;
  }
  return read_exact(p + "_scale", vals + values + meta, meta);
}

} // extern "C"
