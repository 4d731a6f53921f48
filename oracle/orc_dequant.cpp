// see orc_common.h for the header comment (test infrastructure only)
#include "orc_common.h"

// =====================================================================================
// Weight decompression — src/ops/kernels/decompress_kernels.cu:27-70
// =====================================================================================
namespace {
// `DT(q) / scale + offset` with DT = half: operator/ -> __hdiv, operator+ -> __hadd
// (cuda_fp16.hpp:224, 2628, 2723).  Host branch of __hdiv: __float2half(fa / fb); __hadd rounds the
// exact sum once.  dt == 2 applies the same recipe in bf16 (extension; no reference counterpart).
inline uint16_t dq_elem(int q, uint16_t scale_bits, uint16_t off_bits, int dt) {
  if (dt == 1) {
    float fa = (float)q; // __int2half_rn(q) is exact for q <= 255
    float fb = half_bits_to_float(scale_bits);
    uint16_t v = float_to_half_bits(fa / fb);
    double s = (double)half_bits_to_float(v) + (double)half_bits_to_float(off_bits); // exact
    return double_to_half_bits(s);
  }
  float fa = (float)q;
  float fb = bf16_bits_to_float(scale_bits);
  float v = round_bf16(fa / fb);
  return float_to_bf16_bits(v + bf16_bits_to_float(off_bits));
}
} // namespace

extern "C" {

// decompress_int4_general_weights, decompress_kernels.cu:27-53.  valueSize = in_dim*out_dim.
void orc_dequant_int4(uint8_t const *blob, uint16_t *out, int in_dim, int out_dim, int dt) {
  size_t valueSize = (size_t)in_dim * out_dim;
  for (size_t i = 0; i < valueSize / 2; i++) {
    size_t first = i * 2, second = i * 2 + 1;
    size_t group_idx = (first / ((size_t)in_dim * 32)) * in_dim + first % in_dim; // :36-38
    size_t offset_idx = (valueSize / 2) + group_idx * 2;                          // :40
    size_t scale_idx = offset_idx + 2 * (valueSize / 32);                         // :41
    uint16_t sc0, sc1, of0, of1;
    std::memcpy(&sc0, blob + scale_idx, 2);
    std::memcpy(&sc1, blob + scale_idx + 2, 2);
    std::memcpy(&of0, blob + offset_idx, 2);
    std::memcpy(&of1, blob + offset_idx + 2, 2);
    int8_t byte = (int8_t)blob[i]; // char const *, arithmetic shift then & 0xF (:44,48)
    out[first] = dq_elem((byte >> 4) & 0xF, sc0, of0, dt);
    out[second] = dq_elem(byte & 0xF, sc1, of1, dt);
  }
}

// decompress_int8_general_weights, decompress_kernels.cu:55-70.
void orc_dequant_int8(uint8_t const *blob, uint16_t *out, int in_dim, int out_dim, int dt) {
  size_t valueSize = (size_t)in_dim * out_dim;
  for (size_t idx = 0; idx < valueSize; idx++) {
    size_t group_idx = (idx / ((size_t)in_dim * 32)) * in_dim + idx % in_dim;
    size_t offset_idx = valueSize + group_idx * 2;
    size_t scale_idx = offset_idx + 2 * (valueSize / 32);
    uint16_t sc, of;
    std::memcpy(&sc, blob + scale_idx, 2);
    std::memcpy(&of, blob + offset_idx, 2);
    out[idx] = dq_elem(blob[idx] & 0xFF, sc, of, dt);
  }
}

// Quantiser producing the blob above from a [out][in] float matrix.  Rule restated from
// inference/utils/compress_llama_weights.py:15-58 (FlexGen group quantisation: group of 32 along
// dim 0, B = 2^bits-1, scale = B/(mx-mn), q = clamp(round((w-mn)*scale), 0, B)); the blob stores
// offset = mn and scale as DT so that w' = q/scale + mn (decompress_kernels.cu:43-50).  The
// reference's script is truncated before its save step, so the byte packing
// follows the decompress kernel + src/runtime/file_loader.cc:552-669 (even idx = high nibble).
void orc_quantize(float const *w, uint8_t *blob, int in_dim, int out_dim, int bits, int dt) {
  size_t valueSize = (size_t)in_dim * out_dim;
  size_t nvalue_bytes = bits == 4 ? valueSize / 2 : valueSize;
  size_t ngroups = valueSize / 32;
  uint8_t *off_base = blob + nvalue_bytes;
  uint8_t *scale_base = off_base + 2 * ngroups;
  float B = bits == 4 ? 15.f : 255.f;
  if (bits == 4) {
    std::memset(blob, 0, nvalue_bytes);
  }
  for (int g = 0; g < out_dim / 32; g++) {
    for (int i = 0; i < in_dim; i++) {
      float mn = FLT_MAX, mx = -FLT_MAX;
      for (int r = 0; r < 32; r++) {
        float v = w[(size_t)(g * 32 + r) * in_dim + i];
        mn = std::min(mn, v);
        mx = std::max(mx, v);
      }
      float scale = B / std::max(mx - mn, 1e-8f);
      uint16_t sb = dt == 2 ? float_to_bf16_bits(scale) : float_to_half_bits(scale);
      uint16_t ob = dt == 2 ? float_to_bf16_bits(mn) : float_to_half_bits(mn);
      size_t gi = (size_t)g * in_dim + i;
      std::memcpy(scale_base + 2 * gi, &sb, 2);
      std::memcpy(off_base + 2 * gi, &ob, 2);
      for (int r = 0; r < 32; r++) {
        size_t idx = (size_t)(g * 32 + r) * in_dim + i;
        float qf = std::nearbyint((w[idx] - mn) * scale);
        int q = (int)std::min(std::max(qf, 0.f), B);
        if (bits == 4) {
          blob[idx / 2] |= (idx % 2 == 0) ? (uint8_t)(q << 4) : (uint8_t)q;
        } else {
          blob[idx] = (uint8_t)q;
        }
      }
    }
  }
}

} // extern "C"
