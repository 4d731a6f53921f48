// linear_tcc.cu — tcgen05 / TMA projection GEMM for 17..128 tokens with the reduction dimension split over a
// THREAD-BLOCK CLUSTER and summed through distributed shared memory.
// Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel (src/ops/kernels/linear_kernels.cu:507-528) for small
// prompt chunks / small token trees: out[M][N] = A[M][K] * W[N][K]^T, both operands K-major, fp32 accumulation.
//
// Why: at these batch sizes the GEMM is a stream of the weights (7B o_proj: 33.5 MB = 5 us of HBM time), and
// linear_tc's one CTA per 128 output features leaves most of the chip idle when the layer is narrow (o_proj: 32 CTAs on
// 148 SMs, each pulling its 1 MB through one SM's TMA path: 18 us).  Here a cluster of S CTAs (S = 2, 4 or 8, chosen so
// that tiles x S fits the SMs) shares one 128-feature tile: CTA r runs UMMA 128 x 128 x 16 (cta_group::1) over its S-th
// of the k-blocks, parks its fp32 partial tile in its own shared memory (the operand ring is idle by then), and after a
// cluster barrier reduces ITS S-th of the tile's columns, reading all S partials over DSMEM in rank order - a fixed
// order, no workspace in global memory, no counters.  (A stream-K schedule with partials through L2 and a last-arriver
// merge was measured first: bit-exact but slower than linear_tc - 21.9 against 18.1 us at o_proj, 64 tokens - because
// one warp's chain of dependent L2 reads does the merge; profiles/README.md.)
#include <algorithm>
#include <cstdio>
#include <type_traits>

#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int CC_BM = 128, CC_BK = 64, CC_THREADS = 256, CC_MAX_STAGES = 6;
constexpr uint32_t CC_A_BYTES = CC_BM * CC_BK * 2;
constexpr size_t CC_RING_BYTES = (size_t)192 << 10, CC_SMEM = CC_RING_BYTES + 1024; // the partial tile (<= 128 KB) reuses the ring

__device__ __forceinline__ uint32_t cc_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cc_nctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_nctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t cc_mapa(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ float cc_ld_dsmem(uint32_t addr) {
  float v;
  asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void cc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// a_rows: rows of the token box (M rounded up to 8): rows of the A tile past it keep whatever the ring held - they only
// reach accumulator rows >= M, which are never read
template <typename T>
__global__ void __launch_bounds__(CC_THREADS, 1)
linear_tcc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K, int a_rows, int CC_BN,
                  int CC_STAGES) {
  // CC_BN: tile width = UMMA N (128 by default; 192 / 256 by FFB_TCC_BN)
  uint32_t const CC_B_BYTES = (uint32_t)CC_BN * CC_BK * 2, CC_STAGE = CC_A_BYTES + CC_B_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[CC_MAX_STAGES], empty_bar[CC_MAX_STAGES], tmem_full_bar;
  __shared__ uint32_t tmem_base_smem;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u); // 128B-swizzled tiles: 1024-byte alignment

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int const S = (int)cc_nctarank(), rank = (int)cc_ctarank();
  int const tile = (int)blockIdx.x / S, m0 = (int)blockIdx.y * CC_BM; // (clusters run along x: one tile, one 128-token block)
  int const num_k = K / CC_BK;
  int const kb0 = (int)((long)rank * num_k / S), kb1 = (int)((long)(rank + 1) * num_k / S); // (host: num_k >= S)
  uint32_t const stage_tx = (uint32_t)a_rows * CC_BK * 2 + CC_B_BYTES;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < CC_STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  uint32_t const tmem_cols = CC_BN > 128 ? 256u : 128u;
  if (warp == 2) { // CC_BN fp32 accumulator columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    // programmatic dependent launch: the weights do not depend on the preceding kernel - the first ring fill's weight
    // halves go out before the wait, the token halves after it
    int const pre = min(CC_STAGES, kb1 - kb0);
    for (int i = 0; i < pre; i++) {
      mbar_expect_tx(&full_bar[i], stage_tx);
      tma_load_2d(smem + i * CC_STAGE + CC_A_BYTES, &tmB, (kb0 + i) * CC_BK, tile * CC_BN, &full_bar[i]);
    }
    pdl_wait();
    pdl_trigger();
    for (int n = 0; n < kb1 - kb0; n++) {
      int const s = n % CC_STAGES;
      uint8_t *const st = smem + s * CC_STAGE;
      if (n >= pre) {
        mbar_wait(&empty_bar[s], ((n / CC_STAGES) & 1) ^ 1);
        mbar_expect_tx(&full_bar[s], stage_tx);
        tma_load_2d(st + CC_A_BYTES, &tmB, (kb0 + n) * CC_BK, tile * CC_BN, &full_bar[s]);
      }
      tma_load_2d(st, &tmA, (kb0 + n) * CC_BK, m0, &full_bar[s]);
    }
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
    uint32_t const idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(CC_BN >> 3) << 17) | ((uint32_t)(CC_BM >> 4) << 24);
    for (int n = 0; n < kb1 - kb0; n++) {
      int const s = n % CC_STAGES;
      mbar_wait(&full_bar[s], (n / CC_STAGES) & 1);
      tc_fence_after();
      uint32_t const sa = smem_u32(smem + s * CC_STAGE);
      uint64_t const da = umma_desc_sw128(sa), db = umma_desc_sw128(sa + CC_A_BYTES);
#pragma unroll
      for (int k = 0; k < CC_BK / 16; k++) {
        tc_mma_f16(tmem_base, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (uint32_t)((n | k) != 0));
      }
      tc_commit(&empty_bar[s]);
    }
    tc_commit(&tmem_full_bar);
  }
  // ===== epilogue, step 1 (warps 4-7): the partial tile, TMEM -> this CTA's shared memory as [column][row] fp32 =====
  float *const part = reinterpret_cast<float *>(smem);
  int const q = warp & 3, row = q * 32 + lane, row_g = m0 + row;
  bool const epi = warp >= 4, has_rows = m0 + q * 32 < M; // (uniform across the cluster)
  int const col0 = tile * CC_BN;
  if (epi) {
    mbar_wait(&tmem_full_bar, 0); // every UMMA has completed: the ring is idle, the accumulator final
    tc_fence_after();
    if (has_rows) {
      uint32_t const tsrc = tmem_base + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < CC_BN; c0 += 32) {
        uint32_t r[32];
        tmem_ld32(tsrc + (uint32_t)c0, r);
        if (S == 1) { // no split: store
          if (row_g < M) {
#pragma unroll
            for (int v = 0; v < 4; v++) {
              int const col = col0 + c0 + v * 8;
              if (col < N) { // N % 8 == 0
                float f[8];
#pragma unroll
                for (int e = 0; e < 8; e++) {
                  f[e] = __uint_as_float(r[v * 8 + e]) + (bias ? DT<T>::to_f(bias[col + e]) : 0.f);
                }
                *reinterpret_cast<uint4 *>(out + (size_t)row_g * N + col) = pack8<T>(f);
              }
            }
          }
        } else {
#pragma unroll
          for (int e = 0; e < 32; e++) {
            part[(c0 + e) * CC_BM + row] = __uint_as_float(r[e]);
          }
        }
      }
    }
    tc_fence_before();
  }
  if (S > 1) {
    // ===== step 2: every CTA of the cluster sums its S-th of the columns over all partials, in rank order =====
    cc_cluster_sync(); // all partials are in place (release / acquire at cluster scope)
    if (epi && has_rows) {
      // columns [c_begin, c_end): CC_BN / S rounded up to 16 per rank (S = 3: 48 / 48 / 32 of 128), the last ranks take less
      int const cw = (CC_BN / S + 15) / 16 * 16, c_begin = rank * cw, c_end = min(CC_BN, c_begin + cw);
      uint32_t const base = smem_u32(part) + (uint32_t)row * 4u;
      uint32_t src[8]; // (host: S <= 8)
      for (int j = 0; j < S; j++) {
        src[j] = cc_mapa(base, (uint32_t)j);
      }
#pragma unroll 1
      for (int c0 = c_begin; c0 < c_end; c0 += 16) {
        float acc[16];
#pragma unroll
        for (int e = 0; e < 16; e++) {
          acc[e] = 0.f;
        }
        for (int j = 0; j < S; j++) {
#pragma unroll
          for (int e = 0; e < 16; e++) {
            acc[e] += cc_ld_dsmem(src[j] + (uint32_t)((c0 + e) * CC_BM) * 4u);
          }
        }
        if (row_g < M) {
#pragma unroll
          for (int v = 0; v < 2; v++) {
            int const col = col0 + c0 + v * 8;
            if (col < N) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; e++) {
                f[e] = acc[v * 8 + e] + (bias ? DT<T>::to_f(bias[col + e]) : 0.f);
              }
              *reinterpret_cast<uint4 *>(out + (size_t)row_g * N + col) = pack8<T>(f);
            }
          }
        }
      }
    }
    cc_cluster_sync(); // no CTA leaves while a peer may still read its shared memory
  } else {
    __syncthreads();
  }
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(tmem_cols));
  }
}

template <typename T>
int launch_cc(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN, int S,
              cudaStream_t s) {
  int const a_rows = std::min(CC_BM, (M + 7) / 8 * 8); // (more than 128 tokens: full boxes, the last block zero-filled)
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, a_rows);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, BN);
  if (rc != FFB_OK) {
    return rc;
  }
  FFB_SET_MAX_SMEM_ONCE(linear_tcc_kernel<T>, CC_SMEM);
  int const n_tiles = (N + BN - 1) / BN;
  int const stages = std::min<int>(CC_MAX_STAGES, (int)(CC_RING_BYTES / (CC_A_BYTES + (size_t)BN * CC_BK * 2)));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(n_tiles * S, (M + CC_BM - 1) / CC_BM);
  cfg.blockDim = dim3(CC_THREADS);
  cfg.dynamicSmemBytes = CC_SMEM;
  cfg.stream = s;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = S;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = gemm_pdl_enabled() ? 2 : 1;
  FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, linear_tcc_kernel<T>, tmA, tmB, (T *)out, (T const *)bias, M, N, K, a_rows, BN, stages));
  return FFB_OK;
}

} // namespace

// Tile width and k-split.  All clusters must be resident at once: a 148-SM B200 holds 33 clusters of 4 CTAs and 74 of 2
// (GPCs whose SM count is not a multiple of the cluster size lose some), and a cluster beyond that starts when the
// first ones exit (36 clusters of 4: 21.4 us where the SM-pair kernel takes 17.5).  Default: 128-feature tiles, S = 4
// when (token blocks x tiles) <= 33, S = 2 when <= 74, each CTA keeping >= 4 k-blocks, else 1 (7B o_proj up to 128
// tokens: 32 tiles x 4).  Measured alternatives (FFB_TCC_BN = 128 / 192 / 256, FFB_TCC_SPLIT = 1 .. 8 force them): wider
// tiles with a deeper split do not help - qkv_proj at 64 tokens is 22.8 us with 64 tiles of 192 x 2 against 22.0 us with
// 96 tiles of 128 (the layer is at its HBM time plus fixed costs either way); clusters of 8 are far slower (o_proj, 16
// tiles of 256 x 8: 23.9 against 11.7 us); clusters of 3 (48 of them) 25.5 against 17.7 us.
constexpr int kClusters4 = 33, kClusters2 = kNumSMs / 2;

void linear_tcc_pick(int M, int K, int N, int *bn_out, int *s_out) {
  int const num_k = K / CC_BK, units = ((M + CC_BM - 1) / CC_BM) * ((N + 127) / 128);
  int best_bn = 128, best_s = 1;
  if (units <= kClusters4 && num_k >= 16) {
    best_s = 4;
  } else if (units <= kClusters2 && num_k >= 8) {
    best_s = 2;
  }
  if (char const *e = env("FFB_TCC_BN")) {
    int const v = atoi(e);
    if (v == 128 || v == 192 || v == 256) {
      best_bn = v;
    }
  }
  if (char const *e = env("FFB_TCC_SPLIT")) {
    int const v = atoi(e);
    if (v >= 1 && v <= 8 && num_k >= v) {
      best_s = v;
    }
  }
  *bn_out = best_bn;
  *s_out = best_s;
}

// Up to 128 tokens of dense weights: always.  Above that (up to 512 tokens, K >= 1024) only where the layer is so
// narrow that the SM-pair kernel has a handful of tiles and at most its two-way split, and the clusters fill the chip:
// (token blocks x 128-feature tiles) <= 33 (S = 4: 256 x 4096 x 2048 12.5 against 17.3 us, 512 x 4096 x 1024 12.2 against
// 17.4 us) or 56 .. 74 (S = 2: o_proj at 256 tokens 16.1 against 18.5 us).
// FFB_TCC=0 disables it (linear_tc's one CTA per 128 features / linear_tc2 then).
int linear_tcc_supported(int dtype, int M, int K, int N) {
  if (char const *e = env("FFB_TCC")) {
    if (e[0] == '0') {
      return 0;
    }
  }
  if (!((dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && M > 0 && K % CC_BK == 0 && N % 8 == 0)) {
    return 0;
  }
  if (M <= CC_BM) {
    return 1;
  }
  int const units = ((M + CC_BM - 1) / CC_BM) * ((N + 127) / 128);
  return M <= 512 && K >= 1024 && (units <= kClusters4 || (units >= 56 && units <= kClusters2));
}

int launch_linear_tcc(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N, cudaStream_t s) {
  int bn = 128, S = 1;
  linear_tcc_pick(M, K, N, &bn, &S);
  if (dtype == FFB_DT_HALF) {
    return launch_cc<__half>(in, out, w, bias, dtype, M, K, N, bn, S, s);
  }
  return launch_cc<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, S, s);
}

} // namespace ffb
