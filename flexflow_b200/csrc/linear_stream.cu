// linear_stream.cu — weight-streaming projection GEMM for small token counts (decode: M <= 64).
// Replaces Kernels::Linear::forward_kernel (src/ops/kernels/linear_kernels.cu:450-582) for the
// bandwidth-bound case: at M = 16 the cuBLAS call is a stream over the weights, so the kernel is
// built around 128-bit coalesced weight loads (8 warps sweep 512 contiguous bytes of each weight
// row), with the tokens as the M=16 operand of mma.sync m16n8k16 (fp32 accumulate; the reference
// accumulates in DT).  int8/int4 weights (decompress_kernels.cu:27-70 format) are dequantised in
// registers right before the MMA with the bit-exact rule of dequant.cuh — the dense matrix is never
// written to HBM, and the reference's per-forward H2D copy + full decompress pass disappears.
// A block owns NT*8 consecutive output features (inside one 32-row quantisation group, so all its
// warps share scale/offset columns); its 8 warps split K in interleaved chunks and are summed in a
// fixed order through shared memory (deterministic).
#include <cstdlib>

#include <type_traits>

#include "dequant.cuh"

namespace ffb {
namespace {

template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                 uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                        uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

__device__ __forceinline__ uint32_t word(uint4 const &u, int i) {
  return i == 0 ? u.x : (i == 1 ? u.y : (i == 2 ? u.z : u.w));
}
template <typename T> __device__ __forceinline__ uint32_t pack2(T a, T b) {
  return (uint32_t)(*reinterpret_cast<unsigned short *>(&a)) | ((uint32_t)(*reinterpret_cast<unsigned short *>(&b)) << 16);
}

constexpr int LS_WARPS = 8;
// Decode (MT == 1) occupancy, measured on the 7B qkv_proj blob (384 CTAs): int8 with one chunk in flight per warp and
// 3 CTAs per SM (80 registers: every CTA resident in one wave) 23 us, against 35 us with two chunks and 2 CTAs per SM
// (1.3 waves, ncu: long-scoreboard stalls); int4 with 16-byte chunks needs 128 registers (39 us at 3 CTAs per SM, 26.5 us
// at 2), so its decode variant takes 8-byte chunks (int8's register profile, 3 CTAs per SM): 25.4 us - the same time as
// int8 for half the bytes, i.e. both are bound by the per-element instructions of the bit-exact rule, not by HBM.
#ifndef LS_INT4_KE16
#define LS_INT4_KE16 1
#endif
template <int QB> struct LsOcc { static constexpr int BLOCKS = (QB == 8 || (QB == 4 && LS_INT4_KE16)) ? 3 : 2; };

// QB: 0 = DT weights, 8 = int8 blob, 4 = int4 blob.  The k -> mma-slot assignment is a fixed
// permutation applied identically to A and B (a dot product is order-free), chosen so every lane
// reads 16 contiguous bytes of a weight row.
template <typename T, int QB, int NT, int MT>
__global__ void __launch_bounds__(LS_WARPS * 32, MT == 1 ? LsOcc<QB>::BLOCKS : 1)
linear_stream_kernel(T const *__restrict__ A, uint8_t const *__restrict__ W, T const *__restrict__ bias,
                     T *__restrict__ out, int M, int K, int N) {
  // k elements per lane per chunk: 16 B of weights, except int4 decode (MT == 1), which takes 8 B (16 nibbles) so that
  // its register profile is int8's and 3 CTAs fit an SM
  constexpr int KE = QB == 0 ? 8 : (QB == 8 ? 16 : ((MT == 1 && LS_INT4_KE16) ? 16 : 32));
  constexpr int KC = 4 * KE;                            // k elements per warp chunk
  constexpr int STEPS = KE / 4;                         // mma k-steps per chunk
  constexpr int AP = KE / 8;                            // 16-byte activation pieces per row per chunk
  int const n0 = blockIdx.x * (NT * 8);
  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const g = lane >> 2, tq = lane & 3;

  size_t const value_size = (size_t)K * N;
  size_t const nvalue_bytes = QB == 4 ? value_size / 2 : value_size;
  T const *offs = reinterpret_cast<T const *>(W + nvalue_bytes);
  T const *scales = offs + value_size / 32;
  size_t const meta_row = (size_t)(n0 / 32) * K;

  float acc[MT][NT][4];
#pragma unroll
  for (int mt = 0; mt < MT; mt++) {
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      acc[mt][nt][0] = acc[mt][nt][1] = acc[mt][nt][2] = acc[mt][nt][3] = 0.f;
    }
  }

  int const nchunks = K / KC;
  // U chunks per warp are fetched back-to-back BEFORE any of them is consumed, so U*(NT + 2*MT*AP [+ 2*AP])
  // 16-byte loads are in flight per lane: at M <= 16 the kernel is a pure weight stream and its speed is
  // set by bytes in flight, not by the MMAs.
  constexpr int U = MT == 1 ? (QB == 0 ? 4 : 1) : ((MT == 2 && QB == 0) ? 2 : 1);
  constexpr int QP = QB == 0 ? 1 : AP;
  for (int kc0 = warp; kc0 < nchunks; kc0 += LS_WARPS * U) {
    uint4 w[U][NT], sc[U][QP], of[U][QP], alo[U][MT][AP], ahi[U][MT][AP];
#pragma unroll
    for (int u = 0; u < U; u++) {
      int const kc = kc0 + u * LS_WARPS;
      if (kc < nchunks) {
        int const k = kc * KC + tq * KE;
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
          size_t const e = (size_t)(n0 + nt * 8 + g) * K + k;
          if (QB == 4 && KE == 16) {
            uint2 const v = ldg_stream64(W + e / 2);
            w[u][nt] = make_uint4(v.x, v.y, 0, 0);
          } else {
            w[u][nt] = ldg_stream(QB == 0 ? W + e * 2 : (QB == 8 ? W + e : W + e / 2));
          }
        }
        if (QB != 0) {
#pragma unroll
          for (int p = 0; p < QP; p++) {
            sc[u][p] = __ldg(reinterpret_cast<uint4 const *>(scales + meta_row + k + p * 8));
            of[u][p] = __ldg(reinterpret_cast<uint4 const *>(offs + meta_row + k + p * 8));
          }
        }
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
          int const r0 = mt * 16 + g, r1 = r0 + 8;
#pragma unroll
          for (int p = 0; p < AP; p++) {
            alo[u][mt][p] = r0 < M ? __ldg(reinterpret_cast<uint4 const *>(A + (size_t)r0 * K + k + p * 8)) : make_uint4(0, 0, 0, 0);
            ahi[u][mt][p] = r1 < M ? __ldg(reinterpret_cast<uint4 const *>(A + (size_t)r1 * K + k + p * 8)) : make_uint4(0, 0, 0, 0);
          }
        }
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      if (kc0 + u * LS_WARPS < nchunks) {
#pragma unroll
        for (int s = 0; s < STEPS; s++) {
          int const p = (4 * s) / 8, wi = ((4 * s) % 8) / 2; // activation piece / 32-bit word of elements 4s..4s+3
          uint32_t b0[NT], b1[NT];
          if (QB == 0) {
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
              b0[nt] = word(w[u][nt], 2 * s);
              b1[nt] = word(w[u][nt], 2 * s + 1);
            }
          } else if (std::is_same<T, __half>::value && QB != 0) {
            // two-at-a-time bit-exact rule (dequant.cuh): PRMT / LOP3 builds 2^23 + q, one FFMA per element, one
            // packed convert and one HADD2 per pair
            __half const *ps = reinterpret_cast<__half const *>(&sc[u][QB == 0 ? 0 : p]);
            uint32_t const *ow = reinterpret_cast<uint32_t const *>(&of[u][QB == 0 ? 0 : p]);
            int const e0 = (4 * s) % 8;
            DqPairHalf const c01 = dq_pair_coef(ps[e0], ps[e0 + 1], ow[e0 / 2]);
            DqPairHalf const c23 = dq_pair_coef(ps[e0 + 2], ps[e0 + 3], ow[e0 / 2 + 1]);
            if (__all_sync(0xffffffffu, dq_pair_ok(c01) && dq_pair_ok(c23))) {
#pragma unroll
              for (int nt = 0; nt < NT; nt++) {
                uint32_t m0, m1, m2, m3;
                if (QB == 8) {
                  uint32_t const v = word(w[u][nt], s);
                  m0 = __byte_perm(v, kDqMagic, 0x7540);
                  m1 = __byte_perm(v, kDqMagic, 0x7541);
                  m2 = __byte_perm(v, kDqMagic, 0x7542);
                  m3 = __byte_perm(v, kDqMagic, 0x7543);
                } else { // even element = high nibble of its byte (decompress_kernels.cu:43-50)
                  uint32_t const v = word(w[u][nt], s / 2) >> (16 * (s & 1));
                  m0 = ((v >> 4) & 0xFu) | kDqMagic;
                  m1 = (v & 0xFu) | kDqMagic;
                  m2 = ((v >> 12) & 0xFu) | kDqMagic;
                  m3 = ((v >> 8) & 0xFu) | kDqMagic;
                }
                b0[nt] = dq_pair(m0, m1, c01);
                b1[nt] = dq_pair(m2, m3, c23);
              }
            } else { // a scale of 0 / NaN somewhere in the warp's columns: element-wise rule
              DqCoefHalf cf[4];
              __half const *po = reinterpret_cast<__half const *>(&of[u][QB == 0 ? 0 : p]);
#pragma unroll
              for (int e = 0; e < 4; e++) {
                cf[e] = dq_coef(ps[e0 + e], po[e0 + e]);
              }
#pragma unroll
              for (int nt = 0; nt < NT; nt++) {
                int q0, q1, q2, q3;
                if (QB == 8) {
                  uint32_t const v = word(w[u][nt], s);
                  q0 = v & 0xFF; q1 = (v >> 8) & 0xFF; q2 = (v >> 16) & 0xFF; q3 = v >> 24;
                } else {
                  uint32_t const v = (word(w[u][nt], s / 2) >> (16 * (s & 1))) & 0xFFFFu;
                  q0 = (v >> 4) & 0xF; q1 = v & 0xF; q2 = (v >> 12) & 0xF; q3 = (v >> 8) & 0xF;
                }
                b0[nt] = pack2<__half>(dq_elem(q0, cf[0]), dq_elem(q1, cf[1]));
                b1[nt] = pack2<__half>(dq_elem(q2, cf[2]), dq_elem(q3, cf[3]));
              }
            }
          } else {
            typename DqCoef<T>::type cf[4];
            T const *ps = reinterpret_cast<T const *>(&sc[u][QB == 0 ? 0 : p]);
            T const *po = reinterpret_cast<T const *>(&of[u][QB == 0 ? 0 : p]);
#pragma unroll
            for (int e = 0; e < 4; e++) {
              cf[e] = dq_coef(ps[(4 * s) % 8 + e], po[(4 * s) % 8 + e]);
            }
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
              int q0, q1, q2, q3;
              if (QB == 8) {
                uint32_t const v = word(w[u][nt], s);
                q0 = v & 0xFF; q1 = (v >> 8) & 0xFF; q2 = (v >> 16) & 0xFF; q3 = v >> 24;
              } else {
                uint32_t const v = (word(w[u][nt], s / 2) >> (16 * (s & 1))) & 0xFFFFu;
                q0 = (v >> 4) & 0xF; q1 = v & 0xF; q2 = (v >> 12) & 0xF; q3 = (v >> 8) & 0xF;
              }
              b0[nt] = pack2<T>(dq_elem(q0, cf[0]), dq_elem(q1, cf[1]));
              b1[nt] = pack2<T>(dq_elem(q2, cf[2]), dq_elem(q3, cf[3]));
            }
          }
#pragma unroll
          for (int mt = 0; mt < MT; mt++) {
            uint32_t const a0 = word(alo[u][mt][p], wi), a2 = word(alo[u][mt][p], wi + 1);
            uint32_t const a1 = word(ahi[u][mt][p], wi), a3 = word(ahi[u][mt][p], wi + 1);
#pragma unroll
            for (int nt = 0; nt < NT; nt++) {
              mma16816<T>(acc[mt][nt], a0, a1, a2, a3, b0[nt], b1[nt]);
            }
          }
        }
      }
    }
  }

  // fixed-order (deterministic) sum over the 8 k-slices
  __shared__ float tile[MT * 16][NT * 8 + 1];
  if constexpr (MT == 1) {
    // every warp parks its fragment, then each output is summed over the warps in index order
    __shared__ float red[LS_WARPS][16][NT * 8 + 1];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
#pragma unroll
      for (int e = 0; e < 4; e++) {
        red[warp][g + (e >> 1) * 8][nt * 8 + 2 * tq + (e & 1)] = acc[0][nt][e];
      }
    }
    __syncthreads();
    for (int i = tid; i < 16 * NT * 8; i += LS_WARPS * 32) {
      int const r = i / (NT * 8), c = i % (NT * 8);
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < LS_WARPS; w++) {
        v += red[w][r][c];
      }
      tile[r][c] = v;
    }
    __syncthreads();
  } else {
    for (int wsel = 0; wsel < LS_WARPS; wsel++) {
      if (warp == wsel) {
#pragma unroll
        for (int mt = 0; mt < MT; mt++) {
#pragma unroll
          for (int nt = 0; nt < NT; nt++) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
              int const r = mt * 16 + g + (e >> 1) * 8, c = nt * 8 + 2 * tq + (e & 1);
              tile[r][c] = wsel == 0 ? acc[mt][nt][e] : tile[r][c] + acc[mt][nt][e];
            }
          }
        }
      }
      __syncthreads();
    }
  }
  constexpr int SEGS = NT; // 8-column (16 B) segments per row
  if (tid < MT * 16 * SEGS) {
    int const r = tid / SEGS, sgm = tid % SEGS;
    if (r < M) {
      float v[8];
#pragma unroll
      for (int e = 0; e < 8; e++) {
        v[e] = tile[r][sgm * 8 + e];
        if (bias) {
          v[e] += DT<T>::to_f(bias[n0 + sgm * 8 + e]);
        }
      }
      *reinterpret_cast<uint4 *>(out + (size_t)r * N + n0 + sgm * 8) = pack8<T>(v);
    }
  }
}

} // namespace

// M <= 64 rows of `in` starting at row m0.  Returns FFB_OK or an error.
int launch_linear_stream(int dtype, int quant, void const *in, void *out, void const *w, void const *bias, int M,
                         int K, int N, cudaStream_t s) {
  int const KC = quant == FFB_DT_NONE ? 32 : (quant == FFB_DT_INT8 ? 64 : 128);
  FFB_REQUIRE(K % KC == 0 && N % 32 == 0, "linear_stream needs in_dim %% %d == 0 and out_dim %% 32 == 0", KC);
  FFB_REQUIRE(M >= 1 && M <= 64, "linear_stream handles 1..64 rows, got %d", M);
  int const MT = M <= 16 ? 1 : (M <= 32 ? 2 : 4);
  int NT = (N / 32 >= 2 * kNumSMs) ? 4 : 2;
  if (char const *e = ffb::env("FFB_LS_NT")) {
    NT = atoi(e) == 4 ? 4 : 2;
  }
  dim3 grid(N / (NT * 8));
#define FFB_LS(TT, QB, NTv, MTv)                                                                             \
  linear_stream_kernel<TT, QB, NTv, MTv><<<grid, LS_WARPS * 32, 0, s>>>((TT const *)in, (uint8_t const *)w, \
                                                                         (TT const *)bias, (TT *)out, M, K, N)
#define FFB_LS_MT(TT, QB, NTv)                                         \
  do {                                                                 \
    if (MT == 1) FFB_LS(TT, QB, NTv, 1);                               \
    else if (MT == 2) FFB_LS(TT, QB, NTv, 2);                          \
    else FFB_LS(TT, QB, NTv, 4);                                       \
  } while (0)
#define FFB_LS_NT(TT, QB)                                              \
  do {                                                                 \
    if (NT == 4) FFB_LS_MT(TT, QB, 4); else FFB_LS_MT(TT, QB, 2);      \
  } while (0)
#define FFB_LS_Q(TT)                                                   \
  do {                                                                 \
    if (quant == FFB_DT_NONE) FFB_LS_NT(TT, 0);                        \
    else if (quant == FFB_DT_INT8) FFB_LS_NT(TT, 8);                   \
    else FFB_LS_NT(TT, 4);                                             \
  } while (0)
  if (dtype == FFB_DT_HALF) {
    FFB_LS_Q(__half);
  } else {
    FFB_LS_Q(__nv_bfloat16);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace ffb
