// linear_qgemv.cu — persistent weight-streaming projection for decode batches (M <= 16 tokens) over the reference's
// int8 / int4 blobs ([values | offsets | scales], src/ops/kernels/decompress_kernels.cu:27-70), DT = half.
// Replaces, for quantised layers, the H2D copy + decompress pass + cublasGemmEx of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:464-528): the blob is streamed once and dequantised on chip with the bit-exact
// rule of dequant.cuh.
//
// Why a second decode kernel beside linear_stream.cu: that kernel loads straight into registers, so every warp sits on
// its loads (ncu: long-scoreboard stalls, 1.3 waves of 2 CTAs per SM, 1.6 TB/s of blob bytes), and the eight lanes
// that share a weight column each recompute its reciprocal.  Here, as in linear_gemv.cu, one CTA per SM stays
// resident and keeps a ~150 KB mbarrier ring of cp.async.bulk copies in flight (two feature rows per warp, issued at the top of every
// stage, right after the CTA-wide barrier that retires the slot's last readers), so HBM latency is paid once per CTA; and the per-column coefficients {rcp(scale), -rcp * 2^23} are computed ONCE per (group, column) by the
// 512 consumer threads (one column pair each, one stage ahead) and read back from shared memory as broadcasts:
//   * a tile is one quantisation group = 32 output features x K; a stage is 32 features x 1024 k
//     (values: one bulk copy per feature row into a 16-byte padded row; raw offsets and scales: one copy each);
//   * 16 consumer warps each own 64 k of every stage (4 mma k-steps); a lane holds 16 CONSECUTIVE k of four feature
//     rows (one 128-bit / 64-bit shared load per row) under a k-permutation applied identically to the activation
//     fragments, which stay in registers for the whole kernel;
//   * shared-memory wavefronts are what a lane's bytes cost, broadcast or not (a 128-bit load is served per quarter
//     warp), so the staged coefficient is the reciprocal alone (4 B per column; -rcp * 2^23 is one FMUL in the lane)
//     and the value rows are padded so that every wavefront is conflict-free;
//   * a zero or negative scale (rcp = inf, or the sign of a zero quotient) switches the stage to the element-wise rule (block-wide OR folded into the
//     per-stage barrier);
//   * the 16 k-slices of a tile are summed in warp order (deterministic).
#include <cstdlib>

#include "dequant.cuh"
#include "tc_common.cuh"

namespace ffb {
namespace {

constexpr int QG_CONSUMERS = 16, QG_THREADS = QG_CONSUMERS * 32; // 16 warps: 128 registers per thread
constexpr int QG_KS = 1024;                // k per stage
constexpr int QG_RING_BYTES = 160 * 1024;
constexpr int QG_RED_STRIDE = 36;          // floats per token row of a warp's partial tile (32 + pad)

// barrier over the consumer threads that also ORs a predicate across them
__device__ __forceinline__ int qg_bar_or(int mine) {
  int res;
  asm volatile(
      "{\n\t"
      ".reg .pred p, q;\n\t"
      "setp.ne.u32 q, %1, 0;\n\t"
      "bar.red.or.pred p, 1, %2, q;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t"
      "}" : "=r"(res) : "r"(mine), "n"(QG_CONSUMERS * 32) : "memory");
  return res;
}
__device__ __forceinline__ void qg_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                       uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t qg_word(uint4 const &u, int i) {
  return i == 0 ? u.x : (i == 1 ? u.y : (i == 2 ? u.z : u.w));
}
__device__ __forceinline__ uint32_t qg_pack(__half a, __half b) {
  return (uint32_t)__half_as_ushort(a) | ((uint32_t)__half_as_ushort(b) << 16);
}

template <int QB> struct QgCfg {
  static constexpr int ROWB = QG_KS * QB / 8;   // value bytes of one feature row per stage
  // padded so that the lanes served by one shared-memory wavefront hit distinct banks: a 128-bit load is served per
  // quarter warp (rows 2q, 2q+1 x 4 lanes of 16 B: row stride = 64 mod 128), a 64-bit load per half warp (rows
  // 4h .. 4h+3 x 4 lanes of 8 B: row stride = 32 mod 128)
  static constexpr int WROW = ROWB + (QB == 8 ? 64 : 32);
  static constexpr int VAL_BYTES = 32 * WROW;
  static constexpr int META_BYTES = QG_KS * 2;  // raw offsets, then raw scales
  static constexpr int STAGE_BYTES = VAL_BYTES + 2 * META_BYTES;
  static constexpr int STAGES = QG_RING_BYTES / STAGE_BYTES; // 4 (int8) / 7 (int4)
  static constexpr int COEF_BYTES = 2 * QG_KS * 4;            // two buffers of rcp(scale) per column
  static constexpr int RED_BYTES = QG_CONSUMERS * 16 * QG_RED_STRIDE * 4;
  static constexpr size_t SMEM = (size_t)STAGES * STAGE_BYTES + COEF_BYTES + RED_BYTES;
};

struct QgAcc {
  float v[4][4];
};
// One stage with the element-wise rule (a scale of zero: rcp = inf, where the pair rule's fma(inf, m, -inf) would give
// NaN for every q instead of inf for q >= 1).  Kept out of line so that its registers do not weigh on the fast path.
template <int QB>
__device__ __noinline__ QgAcc qg_slow_stage(QgAcc acc, uint4 w0, uint4 w1, uint4 w2, uint4 w3, uint4 al0, uint4 al1,
                                            uint4 ah0, uint4 ah1, uint8_t const *scales_raw, uint8_t const *offs_raw,
                                            int kl) {
  uint4 const wv[4] = {w0, w1, w2, w3};
#pragma unroll
  for (int ks = 0; ks < 4; ks++) {
    DqCoefHalf cf[4];
#pragma unroll
    for (int e = 0; e < 4; e++) {
      int const col = kl + 4 * ks + e;
      cf[e] = dq_coef(*reinterpret_cast<__half const *>(scales_raw + 2 * col),
                      *reinterpret_cast<__half const *>(offs_raw + 2 * col));
    }
    uint4 const &al = ks < 2 ? al0 : al1, &ah = ks < 2 ? ah0 : ah1;
    uint32_t const a0 = qg_word(al, (ks & 1) * 2), a2 = qg_word(al, (ks & 1) * 2 + 1);
    uint32_t const a1 = qg_word(ah, (ks & 1) * 2), a3 = qg_word(ah, (ks & 1) * 2 + 1);
#pragma unroll
    for (int nt = 0; nt < 4; nt++) {
      int q0, q1, q2, q3;
      if (QB == 8) {
        uint32_t const v = qg_word(wv[nt], ks);
        q0 = v & 0xFF; q1 = (v >> 8) & 0xFF; q2 = (v >> 16) & 0xFF; q3 = v >> 24;
      } else {
        uint32_t const v = (qg_word(wv[nt], ks >> 1) >> (16 * (ks & 1))) & 0xFFFFu;
        q0 = (v >> 4) & 0xF; q1 = v & 0xF; q2 = (v >> 12) & 0xF; q3 = (v >> 8) & 0xF;
      }
      qg_mma(acc.v[nt], a0, a1, a2, a3, qg_pack(dq_elem(q0, cf[0]), dq_elem(q1, cf[1])),
             qg_pack(dq_elem(q2, cf[2]), dq_elem(q3, cf[3])));
    }
  }
  return acc;
}

// SPT = K / 1024 stages per tile (1, 2 or 4)
template <int QB, int SPT>
__global__ void __launch_bounds__(QG_THREADS, 1)
linear_qgemv_kernel(__half const *__restrict__ A, uint8_t const *__restrict__ W, __half const *__restrict__ bias,
                    __half *__restrict__ out, int M, int N) {
  using C = QgCfg<QB>;
  constexpr int K = SPT * QG_KS;
  constexpr int STAGES = C::STAGES;
  extern __shared__ __align__(16) uint8_t qg_smem[];
  uint8_t *ring = qg_smem;
  float *coef = reinterpret_cast<float *>(ring + (size_t)STAGES * C::STAGE_BYTES); // [2][1024] rcp(scale)
  float *red = reinterpret_cast<float *>(reinterpret_cast<uint8_t *>(coef) + C::COEF_BYTES);
  __shared__ __align__(8) uint64_t full_bar[STAGES];

  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const ngroups = N / 32;
  int const my_tiles = (ngroups - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
  int const total_it = my_tiles * SPT;

  size_t const value_size = (size_t)K * N;
  size_t const nvalue_bytes = QB == 4 ? value_size / 2 : value_size;
  __half const *offs = reinterpret_cast<__half const *>(W + nvalue_bytes);
  __half const *scales = offs + value_size / 32;

  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      tc::mbar_init(&full_bar[s], 1);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();

  // Stage `j` of this CTA's stream into ring slot j % STAGES: lane 0 of warp w copies feature rows 2w and 2w + 1
  // (a per-lane copy would be serialised by the compiler: cp.async.bulk takes uniform registers), warp 0 also the raw
  // offsets / scales.  No empty barriers: stage j is requested at the top of iteration
  // j - (STAGES - 1), i.e. right after the CTA-wide barrier that ends iteration j - STAGES, the last reader of the slot.
  auto request = [&](int j) {
    if (lane != 0) {
      return;
    }
    int const t = j / SPT, st = j - t * SPT, s = j % STAGES;
    int const gi = (int)blockIdx.x + t * (int)gridDim.x;
    uint8_t *stage = ring + (size_t)s * C::STAGE_BYTES;
    if (warp == 0) { // the transaction count may run negative until this arrives; the phase cannot complete before it
      tc::mbar_expect_tx(&full_bar[s], 32 * C::ROWB + 2 * C::META_BYTES);
      size_t const mrow = (size_t)gi * K + (size_t)st * QG_KS;
      tc::bulk_load_1d(stage + C::VAL_BYTES, offs + mrow, C::META_BYTES, &full_bar[s]);
      tc::bulk_load_1d(stage + C::VAL_BYTES + C::META_BYTES, scales + mrow, C::META_BYTES, &full_bar[s]);
    }
#pragma unroll
    for (int i = 0; i < 2; i++) { // warp w copies feature rows 2w and 2w + 1
      int const r = 2 * warp + i;
      size_t const e = (size_t)(gi * 32 + r) * K + (size_t)st * QG_KS; // element index of the row segment
      tc::bulk_load_1d(stage + r * C::WROW, W + (QB == 4 ? e / 2 : e), C::ROWB, &full_bar[s]);
    }
  };
  for (int j = 0; j < STAGES - 1 && j < total_it; j++) {
    request(j);
  }

  int const g = lane >> 2, tq = lane & 3;
  int const kl = warp * 64 + 16 * tq; // first of this lane's 16 consecutive k inside a stage
  uint4 alo[SPT][2], ahi[SPT][2];     // activations of token rows g and g + 8 at those k, for every stage of a tile
#pragma unroll
  for (int st = 0; st < SPT; st++) {
#pragma unroll
    for (int p = 0; p < 2; p++) {
      int const k = st * QG_KS + kl + 8 * p;
      alo[st][p] = g < M ? __ldg(reinterpret_cast<uint4 const *>(A + (size_t)g * K + k)) : make_uint4(0, 0, 0, 0);
      ahi[st][p] = g + 8 < M ? __ldg(reinterpret_cast<uint4 const *>(A + (size_t)(g + 8) * K + k)) : make_uint4(0, 0, 0, 0);
    }
  }

  // coefficients of stage slot s into buffer b: thread tid owns column pair tid
  auto coef_phase = [&](int s, int b) -> int {
    uint8_t const *stage = ring + (size_t)s * C::STAGE_BYTES;
    uint32_t const sc2 = *reinterpret_cast<uint32_t const *>(stage + C::VAL_BYTES + C::META_BYTES + 4 * tid);
    __half2 const h2 = *reinterpret_cast<__half2 const *>(&sc2);
    float2 c;
    c.x = rcp_approx_ftz(__low2float(h2));
    c.y = rcp_approx_ftz(__high2float(h2));
    *reinterpret_cast<float2 *>(coef + b * QG_KS + 2 * tid) = c;
    return (dq_rcp_ok(c.x) && dq_rcp_ok(c.y)) ? 0 : 1; // inf / NaN / negative: element-wise rule for this stage
  };

  int bad_cur = 0;
  if (total_it > 0) {
    tc::mbar_wait(&full_bar[0], 0);
    bad_cur = qg_bar_or(coef_phase(0, 0));
  }
  for (int t = 0; t < my_tiles; t++) {
    int const n0 = ((int)blockIdx.x + t * (int)gridDim.x) * 32;
    float acc[4][4];
#pragma unroll
    for (int nt = 0; nt < 4; nt++) {
      acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
    }
#pragma unroll
    for (int st = 0; st < SPT; st++) {
      int const it = t * SPT + st, s = it % STAGES;
      int bad_next = 0;
      if (it + STAGES - 1 < total_it) {
        request(it + STAGES - 1);
      }
      if (it + 1 < total_it) { // next stage's coefficients first: its reciprocals are ready when this stage is done
        int const s1 = (it + 1) % STAGES;
        tc::mbar_wait(&full_bar[s1], ((it + 1) / STAGES) & 1);
        bad_next = coef_phase(s1, (it + 1) & 1);
      }
      uint8_t const *stage = ring + (size_t)s * C::STAGE_BYTES;
      uint8_t const *offs_raw = stage + C::VAL_BYTES, *scales_raw = offs_raw + C::META_BYTES;
      uint4 wv[4]; // int8: 16 bytes, int4: 8 bytes (in .x/.y) of feature rows g, g+8, g+16, g+24
#pragma unroll
      for (int nt = 0; nt < 4; nt++) {
        uint8_t const *rowp = stage + (nt * 8 + g) * C::WROW;
        if (QB == 8) {
          wv[nt] = *reinterpret_cast<uint4 const *>(rowp + kl);
        } else {
          uint2 const v = *reinterpret_cast<uint2 const *>(rowp + kl / 2);
          wv[nt] = make_uint4(v.x, v.y, 0, 0);
        }
      }
      uint4 const o01 = *reinterpret_cast<uint4 const *>(offs_raw + 2 * kl);
      uint4 const o23 = *reinterpret_cast<uint4 const *>(offs_raw + 2 * kl + 16);
      if (!bad_cur) {
        float4 const *cb = reinterpret_cast<float4 const *>(coef + (it & 1) * QG_KS + kl);
#pragma unroll
        for (int ks = 0; ks < 4; ks++) { // elements 4ks .. 4ks+3 of the lane's 16
          float4 const r4 = cb[ks];
          DqPairHalf c01, c23;
          c01.r0 = r4.x; c01.c0 = -r4.x * 8388608.f; c01.r1 = r4.y; c01.c1 = -r4.y * 8388608.f;
          c23.r0 = r4.z; c23.c0 = -r4.z * 8388608.f; c23.r1 = r4.w; c23.c1 = -r4.w * 8388608.f;
          c01.off2 = qg_word(ks < 2 ? o01 : o23, (2 * ks) & 3);
          c23.off2 = qg_word(ks < 2 ? o01 : o23, (2 * ks + 1) & 3);
          uint32_t const a0 = qg_word(alo[st][ks >> 1], (ks & 1) * 2), a2 = qg_word(alo[st][ks >> 1], (ks & 1) * 2 + 1);
          uint32_t const a1 = qg_word(ahi[st][ks >> 1], (ks & 1) * 2), a3 = qg_word(ahi[st][ks >> 1], (ks & 1) * 2 + 1);
#pragma unroll
          for (int nt = 0; nt < 4; nt++) {
            uint32_t m0, m1, m2, m3;
            if (QB == 8) {
              uint32_t const v = qg_word(wv[nt], ks);
              m0 = __byte_perm(v, kDqMagic, 0x7540);
              m1 = __byte_perm(v, kDqMagic, 0x7541);
              m2 = __byte_perm(v, kDqMagic, 0x7542);
              m3 = __byte_perm(v, kDqMagic, 0x7543);
            } else { // even element = high nibble of its byte (decompress_kernels.cu:43-50)
              uint32_t const v = qg_word(wv[nt], ks >> 1) >> (16 * (ks & 1));
              m0 = ((v >> 4) & 0xFu) | kDqMagic;
              m1 = (v & 0xFu) | kDqMagic;
              m2 = ((v >> 12) & 0xFu) | kDqMagic;
              m3 = ((v >> 8) & 0xFu) | kDqMagic;
            }
#if defined(QG_EXP_NO_MMA) // experiment builds (tools/build_variant.sh): where does the time go
            acc[nt][0] += __uint_as_float(dq_pair(m0, m1, c01) ^ a0);
            acc[nt][1] += __uint_as_float(dq_pair(m2, m3, c23) ^ a1);
#elif defined(QG_EXP_NO_DQ)
            qg_mma(acc[nt], a0, a1, a2, a3, m0 ^ m1, m2 ^ m3);
#else
            qg_mma(acc[nt], a0, a1, a2, a3, dq_pair(m0, m1, c01), dq_pair(m2, m3, c23));
#endif
          }
        }
      } else { // a zero scale somewhere in this stage: element-wise rule with the raw scales (out of line)
        QgAcc r;
#pragma unroll
        for (int nt = 0; nt < 4; nt++) {
#pragma unroll
          for (int e = 0; e < 4; e++) {
            r.v[nt][e] = acc[nt][e];
          }
        }
        r = qg_slow_stage<QB>(r, wv[0], wv[1], wv[2], wv[3], alo[st][0], alo[st][1], ahi[st][0], ahi[st][1], scales_raw,
                              offs_raw, kl);
#pragma unroll
        for (int nt = 0; nt < 4; nt++) {
#pragma unroll
          for (int e = 0; e < 4; e++) {
            acc[nt][e] = r.v[nt][e];
          }
        }
      }
      if (st == SPT - 1) { // park this warp's k-slice of the tile
        float *slot = red + (size_t)warp * 16 * QG_RED_STRIDE;
#pragma unroll
        for (int nt = 0; nt < 4; nt++) {
          *reinterpret_cast<float2 *>(slot + g * QG_RED_STRIDE + nt * 8 + 2 * tq) = make_float2(acc[nt][0], acc[nt][1]);
          *reinterpret_cast<float2 *>(slot + (g + 8) * QG_RED_STRIDE + nt * 8 + 2 * tq) = make_float2(acc[nt][2], acc[nt][3]);
        }
      }
      bad_cur = qg_bar_or(bad_next);
      if (st == SPT - 1) { // sum the 16 k-slices in warp order: thread -> (token tid / 32, feature tid % 32)
        int const r = tid >> 5, c = tid & 31;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < QG_CONSUMERS; w++) {
          v += red[((size_t)w * 16 + r) * QG_RED_STRIDE + c];
        }
        if (r < M) {
          if (bias) {
            v += __half2float(bias[n0 + c]);
          }
          out[(size_t)r * N + n0 + c] = __float2half_rn(v);
        }
        if (SPT == 1) { // the next tile parks its partials one stage later: keep them off the sums being read
          asm volatile("bar.sync 1, %0;" ::"n"(QG_CONSUMERS * 32) : "memory");
        }
      }
    }
  }
}

} // namespace

int linear_qgemv_supported(int dtype, int quant, int M, int K, int N) {
  if (ffb::env("FFB_DISABLE_QGEMV")) {
    return 0;
  }
  return dtype == FFB_DT_HALF && (quant == FFB_DT_INT8 || quant == FFB_DT_INT4) && M >= 1 && M <= 16 &&
         (K == 1024 || K == 2048 || K == 4096) && N % 32 == 0;
}

template <int QB, int SPT>
static int launch_qgemv_t(void const *in, void *out, void const *w, void const *bias, int M, int N, cudaStream_t s) {
  constexpr size_t smem = QgCfg<QB>::SMEM;
  FFB_SET_MAX_SMEM_ONCE((linear_qgemv_kernel<QB, SPT>), smem);
  int const ngroups = N / 32;
  int const grid = ngroups < kNumSMs ? ngroups : kNumSMs;
  linear_qgemv_kernel<QB, SPT><<<grid, QG_THREADS, smem, s>>>((__half const *)in, (uint8_t const *)w, (__half const *)bias,
                                                              (__half *)out, M, N);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

int launch_linear_qgemv(int quant, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                        cudaStream_t s) {
  int const spt = K / QG_KS;
  if (quant == FFB_DT_INT8) {
    return spt == 1 ? launch_qgemv_t<8, 1>(in, out, w, bias, M, N, s)
                    : (spt == 2 ? launch_qgemv_t<8, 2>(in, out, w, bias, M, N, s) : launch_qgemv_t<8, 4>(in, out, w, bias, M, N, s));
  }
  return spt == 1 ? launch_qgemv_t<4, 1>(in, out, w, bias, M, N, s)
                  : (spt == 2 ? launch_qgemv_t<4, 2>(in, out, w, bias, M, N, s) : launch_qgemv_t<4, 4>(in, out, w, bias, M, N, s));
}

} // namespace ffb
