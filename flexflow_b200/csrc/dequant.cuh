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
