// linear_tcq.cu — tcgen05 / TMA projection GEMM over QUANTISED weights for batched prompt / tree tokens (M > 64):
// the int8 / int4 values are dequantised IN SHARED MEMORY, tile by tile, right before the MMA (north_star kernel family
// 1).  Replaces, for offloaded / quantised layers, the reference's decompress-the-whole-matrix-to-HBM + cublasGemmEx
// sequence (src/ops/kernels/linear_kernels.cu:461-528) - which linear_tc.cu used to restate with a K x N x 2-byte
// scratch round trip per call.
//
// One 256 x 128 output tile per CTA (two 128-row accumulators in tensor memory sharing every W tile), k-blocks of 64:
//   warp 0 (one lane)   TMA producer: the 256 x 64 block of A (two 128B-swizzled boxes) through a 4-stage mbarrier ring
//   warps 4-11          converters.  Warp w owns output features [32 (w % 4), +32) of the tile - exactly ONE
//                       quantisation group (32 consecutive output rows, src/ops/kernels/decompress_kernels.cu:34-40) -
//                       and k in [32 (w / 4 - 1), +32).  Per k-block each lane reads its feature row's 32 codes (32 or
//                       16 bytes, straight from L2: no staging copy), the warp computes the 32 per-column reciprocals
//                       of its group once (one rcp per lane) and shares them through a 512-byte shared-memory table,
//                       every element takes the bit-exact pair rule of dequant.cuh (magic-number int -> float inside
//                       one FFMA, cvt.rn.f16x2, HADD2 with the offsets; element-wise __hdiv recipe when a reciprocal
//                       is not finite / negative), and the fp16 result is written as the 128B-swizzled K-major B
//                       operand into one of two 16 KB buffers (fence.proxy.async, mbarrier arrive)
//   warp 1 (one lane)   MMA issuer: 2 accumulators x 4 UMMA 128 x 128 x 16 per k-block; tcgen05.commit frees the A
//                       stage and the converted buffer
//   warps 4-11          epilogue after the loop: warps 4-7 rows 0-127, warps 8-11 rows 128-255
// Codes and group meta are read once per 256-row M tile; nothing dequantised ever touches HBM.  fp32 accumulation.
#include <cuda.h>

#include <type_traits>

#include "dequant.cuh"
#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int Q_BM = 256, Q_BN = 128, Q_BK = 64, Q_STAGES = 4, Q_THREADS = 384, Q_CONV_WARPS = 8;
constexpr uint32_t QA_BYTES = Q_BM * Q_BK * 2; // one A stage: two [128][64] fp16 boxes
constexpr uint32_t QB_BYTES = Q_BN * Q_BK * 2; // one converted W tile
constexpr size_t Q_SMEM = (size_t)Q_STAGES * QA_BYTES + 2 * QB_BYTES + 1024;

struct QCoef { // per column of a quantisation group: rcp.approx.ftz(scale), -rcp * 2^23
  float r, c;
};

// QB = 8 or 4 bits per code
template <int QB>
__global__ void __launch_bounds__(Q_THREADS, 1)
linear_tcq_kernel(const __grid_constant__ CUtensorMap tmA, uint8_t const *__restrict__ blob, __half *__restrict__ out,
                  __half const *__restrict__ bias, int M, int N, int K) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[Q_STAGES], empty_bar[Q_STAGES], conv_bar[2], bfree_bar[2], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  __shared__ __align__(16) QCoef s_coef[Q_CONV_WARPS][32];
  __shared__ __align__(16) __half s_off[Q_CONV_WARPS][32];
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + Q_STAGES * QA_BYTES;

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int const n_blk = blockIdx.x, m_blk = blockIdx.y;
  int const num_k = K / Q_BK;
  // blob = [values | offsets | scales] (decompress_kernels.cu:27-70): element (o, i) has code index o * K + i,
  // its group meta index (o / 32) * K + i
  size_t const nvalues = (size_t)N * K;
  uint8_t const *values = blob;
  __half const *offsets = reinterpret_cast<__half const *>(blob + (QB == 4 ? nvalues / 2 : nvalues));
  __half const *scales = offsets + nvalues / 32;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < Q_STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(&conv_bar[b], Q_CONV_WARPS);
      mbar_init(&bfree_bar[b], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) { // two 128-column fp32 accumulators
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(256u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer: A =====
    for (int kb = 0; kb < num_k; kb++) {
      int const s = kb % Q_STAGES;
      mbar_wait(&empty_bar[s], ((kb / Q_STAGES) & 1) ^ 1);
      mbar_expect_tx(&full_bar[s], QA_BYTES);
      tma_load_2d(sA + s * QA_BYTES, &tmA, kb * Q_BK, m_blk * Q_BM, &full_bar[s]);
      tma_load_2d(sA + s * QA_BYTES + QA_BYTES / 2, &tmA, kb * Q_BK, m_blk * Q_BM + 128, &full_bar[s]);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = (1u << 4) | ((uint32_t)(Q_BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24); // F16 x F16 -> F32
    for (int kb = 0; kb < num_k; kb++) {
      int const s = kb % Q_STAGES, b = kb & 1;
      mbar_wait(&full_bar[s], (kb / Q_STAGES) & 1);
      mbar_wait(&conv_bar[b], (kb >> 1) & 1);
      tc_fence_after();
      uint64_t const da0 = umma_desc_sw128(smem_u32(sA + s * QA_BYTES));
      uint64_t const da1 = umma_desc_sw128(smem_u32(sA + s * QA_BYTES + QA_BYTES / 2));
      uint64_t const db = umma_desc_sw128(smem_u32(sB + b * QB_BYTES));
#pragma unroll
      for (int k = 0; k < Q_BK / 16; k++) {
        tc_mma_f16(tmem_base, da0 + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
        tc_mma_f16(tmem_base + 128u, da1 + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
      }
      tc_commit(&empty_bar[s]);
      tc_commit(&bfree_bar[b]);
    }
    tc_commit(&tmem_full_bar);
  } else if (warp >= 4) {
    // ===== converters =====
    int const cw = warp - 4, g = cw & 3, kh = cw >> 2;
    int const n_local = g * 32 + lane;
    size_t const o = (size_t)n_blk * Q_BN + n_local; // this lane's output feature
    size_t const grp = o >> 5;                         // its quantisation group (the same for the whole warp)
    uint8_t const *const qrow = values + (QB == 4 ? (o * (size_t)K) / 2 : o * (size_t)K);
    __half const *const srow = scales + grp * (size_t)K, *const orow = offsets + grp * (size_t)K;
    uint8_t *const brow = sB + (size_t)n_local * 128;
    __shared__ float s_scale[Q_CONV_WARPS][32];
    // next k-block's codes and this lane's column of the group meta, one k-block ahead
    uint4 nq0 = make_uint4(0, 0, 0, 0), nq1 = nq0;
    __half nsc, nof;
    auto fetch = [&](int kb) {
      int const k0 = kb * Q_BK + kh * 32;
      if (QB == 8) {
        nq0 = ldg_stream(qrow + k0);
        nq1 = ldg_stream(qrow + k0 + 16);
      } else {
        nq0 = ldg_stream(qrow + k0 / 2);
      }
      nsc = srow[k0 + lane];
      nof = orow[k0 + lane];
    };
    fetch(0);
    for (int kb = 0; kb < num_k; kb++) {
      int const b = kb & 1;
      uint4 const q0 = nq0, q1 = nq1;
      float const sc = __half2float(nsc);
      float const r = rcp_approx_ftz(sc);
      bool const all_ok = __all_sync(0xffffffffu, dq_rcp_ok(r));
      s_coef[cw][lane] = QCoef{r, -r * 8388608.f};
      s_off[cw][lane] = nof;
      s_scale[cw][lane] = sc;
      if (kb + 1 < num_k) {
        fetch(kb + 1);
      }
      __syncwarp();
      mbar_wait(&bfree_bar[b], ((kb >> 1) & 1) ^ 1); // the MMAs that read this buffer two k-blocks ago have completed
      // codes of this lane's 32 k-values as bytes c[0..31] (int4: element 2j in the HIGH nibble of byte j)
      uint32_t const w8[8] = {q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, q1.z, q1.w};
      auto code = [&](int j) -> uint32_t {
        if (QB == 8) {
          return (w8[j >> 2] >> ((j & 3) * 8)) & 0xFFu;
        }
        uint32_t const byte = (w8[j >> 3] >> (((j >> 1) & 3) * 8)) & 0xFFu;
        return (j & 1) ? (byte & 0xFu) : (byte >> 4);
      };
#pragma unroll
      for (int c = 0; c < 4; c++) {
        uint32_t pk[4];
#pragma unroll
        for (int p = 0; p < 4; p++) {
          int const j = c * 8 + 2 * p;
          float4 const cf = *reinterpret_cast<float4 const *>(&s_coef[cw][j]); // {r_j, c_j, r_j+1, c_j+1}
          uint32_t const off2 = *reinterpret_cast<uint32_t const *>(&s_off[cw][j]);
          if (all_ok) {
            DqPairHalf cp;
            cp.r0 = cf.x; cp.c0 = cf.y; cp.r1 = cf.z; cp.c1 = cf.w; cp.off2 = off2;
            pk[p] = dq_pair(kDqMagic | code(j), kDqMagic | code(j + 1), cp);
          } else { // degenerate scales (0, inf, NaN, negative): the element-wise __hdiv recipe
            __half2 const of2 = *reinterpret_cast<__half2 const *>(&off2);
            __half const e0 = dq_elem((int)code(j), dq_coef(__float2half(s_scale[cw][j]), __low2half(of2)));
            __half const e1 = dq_elem((int)code(j + 1), dq_coef(__float2half(s_scale[cw][j + 1]), __high2half(of2)));
            __half2 const e = __halves2half2(e0, e1);
            pk[p] = *reinterpret_cast<uint32_t const *>(&e);
          }
        }
        int const cc = kh * 4 + c; // 16-byte chunk of the 128-byte operand row, XOR-swizzled by the row (SW128)
        *reinterpret_cast<uint4 *>(brow + (size_t)b * QB_BYTES + ((cc ^ (n_local & 7)) << 4)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      }
      fence_proxy_async(); // generic-proxy writes -> visible to the tensor core's operand reads
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&conv_bar[b]);
      }
    }
    // ===== epilogue: TMEM -> registers -> global =====
    mbar_wait(&tmem_full_bar, 0);
    tc_fence_after();
    int const acc = cw >> 2, q = cw & 3; // accumulator (row half of the tile), TMEM lane quarter
    int const row = m_blk * Q_BM + acc * 128 + q * 32 + lane;
#pragma unroll 1
    for (int c0 = 0; c0 < Q_BN; c0 += 32) {
      uint32_t rr[32];
      tmem_ld32(tmem_base + (uint32_t)(acc * 128) + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, rr);
      if (row < M) {
        int const col = n_blk * Q_BN + c0;
        __half *dst = out + (size_t)row * N + col;
#pragma unroll
        for (int v = 0; v < 4; v++) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; e++) {
            f[e] = __uint_as_float(rr[v * 8 + e]);
            if (bias) {
              f[e] += __half2float(bias[col + v * 8 + e]);
            }
          }
          *reinterpret_cast<uint4 *>(dst + v * 8) = pack8<__half>(f);
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256u));
  }
}

} // namespace

// half activations, int8 / int4 blobs, N a multiple of the 128-wide tile, K of the 64-wide k-block.
// OFF unless FFB_TCQ=1: measured at the 7B shapes (tools/bench_quant.py --m 128 / 512, profiles/README.md) the kernel is
// bound by its converter warps (every 256-row M tile dequantises the whole matrix again, each lane on a dependent
// chain of L2 reads) and loses to dequantise-to-scratch + the SM-pair GEMM: 153 against 89 us (qkv_proj, int8, 512
// tokens), 79 against 41 us (o_proj), 76 against 70 us at 128 tokens.  Kept for the bit-exact test and as the base of
// a version with TMA-staged codes.
int linear_tcq_supported(int dtype, int quant, int K, int N) {
  char const *e = env("FFB_TCQ");
  if (!(e && e[0] == '1') || env("FFB_DISABLE_TCQ")) {
    return 0;
  }
  return dtype == FFB_DT_HALF && (quant == FFB_DT_INT8 || quant == FFB_DT_INT4) && K % Q_BK == 0 && N % Q_BN == 0;
}

int launch_linear_tcq(int quant, void const *in, void *out, void const *blob, void const *bias, int M, int K, int N,
                      cudaStream_t s) {
  CUtensorMap tmA;
  int rc = make_tensor_map_2d(&tmA, FFB_DT_HALF, in, M, K, 128);
  if (rc != FFB_OK) {
    return rc;
  }
  dim3 grid(N / Q_BN, (M + Q_BM - 1) / Q_BM);
  if (quant == FFB_DT_INT8) {
    FFB_SET_MAX_SMEM_ONCE(linear_tcq_kernel<8>, Q_SMEM);
    linear_tcq_kernel<8><<<grid, Q_THREADS, Q_SMEM, s>>>(tmA, (uint8_t const *)blob, (__half *)out, (__half const *)bias, M, N, K);
  } else {
    FFB_SET_MAX_SMEM_ONCE(linear_tcq_kernel<4>, Q_SMEM);
    linear_tcq_kernel<4><<<grid, Q_THREADS, Q_SMEM, s>>>(tmA, (uint8_t const *)blob, (__half *)out, (__half const *)bias, M, N, K);
  }
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace ffb
