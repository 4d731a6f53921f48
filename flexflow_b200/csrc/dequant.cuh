// dequant.cuh — element rule of FlexFlow's int4/int8 group quantisation, shared by the standalone
// decompress kernels and the weight-streaming GEMM so both produce identical bits.
// Reference: src/ops/kernels/decompress_kernels.cu:27-70:  w = DT(q) / scale + offset  in DT arithmetic.
// For DT = half, `/` is __hdiv (cuda_fp16.hpp:2723-2746): rcp.approx.ftz.f32 of the divisor, one fp32
// multiply, round to half, and a fused-multiply-add correction only when the quotient lands in
// (0, 0x008F) (half subnormal range); `+` is __hadd (one rounding).  The reciprocal depends only on
// the scale, so it is computed once per (group, column) and reused for every element of the group:
// the sequence of roundings is unchanged, hence the result is bit-identical to the reference kernel.
#pragma once
#include "common.cuh"

namespace ffb {

__device__ __forceinline__ float rcp_approx_ftz(float x) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
}

struct DqCoefHalf {
  float rcp;   // rcp.approx.ftz(scale)
  float scale; // divisor as float (for the correction step)
  __half off;
};

__device__ __forceinline__ DqCoefHalf dq_coef(__half scale, __half off) {
  DqCoefHalf c;
  c.scale = __half2float(scale);
  c.rcp = rcp_approx_ftz(c.scale);
  c.off = off;
  return c;
}

// DT(q)/scale + offset for DT = half; q in [0, 255]
__device__ __forceinline__ __half dq_elem(int q, DqCoefHalf const &c) {
  float const fa = (float)q; // == __half2float(__int2half_rn(q)), exact
  float fv = c.rcp * fa;
  __half v = __float2half(fv);
  __half const a = __habs(v);
  __half den;
  *reinterpret_cast<unsigned short *>(&den) = 0x008FU;
  if (__hlt(a, den) && __hlt(__float2half(0.0f), a)) {
    float const err = __fmaf_rn(-c.scale, fv, fa);
    fv = __fmaf_rn(c.rcp, err, fv);
    v = __float2half(fv);
  }
  return __hadd(v, c.off);
}

// ---- two elements at a time, no conversion-pipe work per element (DT = half) -----------------------------------------
// dq_elem costs I2F + FMUL + F2H + three half compares + HADD per element, and the I2F / F2H conversions run at a
// quarter of the FP32 rate: the weight-streaming GEMM was conversion-bound at 1.1 TB/s (int8) / 0.6 TB/s (int4).
// Same roundings, cheaper instructions:
//   * float(q): m = 0x4B000000 | q is the float 2^23 + q, and fma(rcp, m, -rcp * 2^23) rounds the exact product
//     rcp * q once - the value FMUL(rcp, float(q)) gives (-rcp * 2^23 is exact: a power-of-two scaling).  Needs a
//     finite, non-negative rcp; dq_pair_ok() says so and the callers fall back to dq_elem otherwise (scale = 0, NaN,
//     negative scales: sign of a zero quotient).
//   * one cvt.rn.f16x2.f32 rounds both quotients, one HADD2 adds both offsets (adjacent in the blob).
//   * __hdiv's correction step runs only for 0 < |quotient| < 0x008F (8.5e-6).  With q >= 1 that needs
//     |rcp| < 8.5e-6, i.e. |scale| > 117000: no half is that large (max 65504; scale = inf gives rcp = 0 and the
//     quotient 0, also outside the range), so the branch is dead for every possible blob.
//     tests/test_gpu_linear.py::test_dequant_exhaustive_quotients checks every q x every finite non-zero half scale
//     through this path against the CPU checker's IEEE division.
struct DqPairHalf {
  float r0, c0, r1, c1; // rcp.approx.ftz(scale) and -rcp * 2^23 of two neighbouring columns
  uint32_t off2;        // their offsets as stored (half2)
};
// Both reciprocals finite AND non-negative (one unsigned compare each: a set sign bit is above 0x7F800000 too).
// Negative scales take the element-wise rule: with q = 0 the reference's product rcp * 0 is -0 for a negative rcp,
// while fma(rcp, 2^23, -rcp * 2^23) cancels exactly to +0, and the two differ after the add when the offset is -0
// (found by running the reference's own kernel on the same blobs, tests/test_gpu_linear.py; compress() never
// produces a negative scale, so real checkpoints stay on the pair rule).
__device__ __forceinline__ bool dq_rcp_ok(float r) { return __float_as_uint(r) < 0x7F800000u; }
__device__ __forceinline__ bool dq_pair_ok(DqPairHalf const &c) { return dq_rcp_ok(c.r0) && dq_rcp_ok(c.r1); }
__device__ __forceinline__ DqPairHalf dq_pair_coef(__half s0, __half s1, uint32_t off2) {
  DqPairHalf c;
  c.r0 = rcp_approx_ftz(__half2float(s0));
  c.r1 = rcp_approx_ftz(__half2float(s1));
  c.c0 = -c.r0 * 8388608.f;
  c.c1 = -c.r1 * 8388608.f;
  c.off2 = off2;
  return c;
}
// m0, m1 = 0x4B000000 | q of the two elements; returns the two dequantised halfs packed (element 0 in the low half)
__device__ __forceinline__ uint32_t dq_pair(uint32_t m0, uint32_t m1, DqPairHalf const &c) {
  float const f0 = __fmaf_rn(c.r0, __uint_as_float(m0), c.c0);
  float const f1 = __fmaf_rn(c.r1, __uint_as_float(m1), c.c1);
  __half2 const v = __floats2half2_rn(f0, f1);
  __half2 const r = __hadd2(v, *reinterpret_cast<__half2 const *>(&c.off2));
  return *reinterpret_cast<uint32_t const *>(&r);
}
constexpr uint32_t kDqMagic = 0x4B000000u;

// bf16 variant (extension — the reference has no bf16): same recipe with bf16 roundings.
struct DqCoefBf16 {
  float rcp;
  float off;
};
__device__ __forceinline__ DqCoefBf16 dq_coef(__nv_bfloat16 scale, __nv_bfloat16 off) {
  DqCoefBf16 c;
  c.rcp = rcp_approx_ftz(__bfloat162float(scale));
  c.off = __bfloat162float(off);
  return c;
}
__device__ __forceinline__ __nv_bfloat16 dq_elem(int q, DqCoefBf16 const &c) {
  float const v = __bfloat162float(__float2bfloat16_rn(c.rcp * (float)q));
  return __float2bfloat16_rn(v + c.off);
}

template <typename T> struct DqCoef;
template <> struct DqCoef<__half> { using type = DqCoefHalf; };
template <> struct DqCoef<__nv_bfloat16> { using type = DqCoefBf16; };

} // namespace ffb
