// linear_w512.cu — the SM-pair (cta_group::2) tcgen05 / TMA projection GEMM with ONE WAVE of wide output tiles
// (256 rows x BN columns, 272 <= BN <= 512).  Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:507-528) for batched prompt / tree tokens where it applies:
// out[M][N] = A[M][K] * W[N][K]^T, both operands K-major, fp32 accumulation.
//
// Why: what a tile costs per k-block is the operand bytes each SM has to take in - (128 + BN / 2) rows of 128 bytes per
// CTA at ~80 GB/s per SM, measured the same for 256 x 256 tiles on all pairs, for 256 x 512 tiles on 48 pairs and for
// clusters that multicast A (profiles/README.md) - and whole waves of tiles.  M = 512, N = 12288 is 96 tiles of
// 256 x 256 on 74 pairs: two waves (linear_tc2), or 1.3 waves' worth plus partial sums through L2 (linear_sk).  Here the
// tile width is chosen so that ALL tiles run in one wave: BN = 336 gives 2 x 37 = 74 tiles, each pair streams
// 64 x (256 + 336) operand rows instead of 83 x 512.  The leader issues TWO UMMAs per k-step against the same A tile
// (column ranges [0, H1) and [H1, BN) of the tile, H1 a multiple of 32, both widths multiples of 16); the accumulator
// takes up to all 512 tensor-memory columns of both CTAs and is single-buffered, which costs nothing with one tile per
// pair.  The last column block may be ragged (TMA zero-fills rows of W past N; the epilogue stores columns < N only).
#include <algorithm>
#include <cstdio>
#include <type_traits>

#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int W5_BM = 128, W5_BK = 64, W5_THREADS = 384, W5_MAX_STAGES = 8; // warps 0-3: TMA / MMA / TMEM alloc / idle, 4-11: epilogue
constexpr uint32_t W5_A_BYTES = W5_BM * W5_BK * 2; // this CTA's 128 rows of A
constexpr size_t W5_RING_BYTES = (size_t)192 << 10, W5_STAGING_BYTES = 8 * 4096; // 8 epilogue warps x 4 KB

__device__ __forceinline__ uint32_t w5_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t w5_mapa(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
__device__ __forceinline__ void w5_tma_load(uint32_t dst, CUtensorMap const *tm, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void w5_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void w5_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(W5_THREADS, 1)
linear_w512_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                   const __grid_constant__ CUtensorMap tmB2, T *__restrict__ out, T const *__restrict__ bias, int M, int N,
                   int K, int BN, int H1, int W5_STAGES) {
  // this CTA's share of W per k-block: H1 / 2 rows for the first UMMA's columns, (BN - H1) / 2 for the second's
  int const H2 = BN - H1;
  uint32_t const B1_BYTES = (uint32_t)(H1 / 2) * W5_BK * 2, B2_BYTES = (uint32_t)(H2 / 2) * W5_BK * 2; // multiples of 1024
  uint32_t const W5_STAGE = W5_A_BYTES + B1_BYTES + B2_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[W5_MAX_STAGES], empty_bar[W5_MAX_STAGES], tmem_full_bar, tmem_empty_bar;
  __shared__ uint32_t tmem_base_smem;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *stage_out = smem + W5_STAGES * W5_STAGE; // 8 epilogue warps x 4 KB

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = w5_ctarank(); // 0 = leader
  int const num_k = K / W5_BK;
  int const m_tiles = (M + 255) / 256, n_tiles = (N + BN - 1) / BN, tiles = m_tiles * n_tiles;
  int const pair = blockIdx.x >> 1, P = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB2) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < W5_STAGES; s++) {
      mbar_init(&full_bar[s], 1);  // leader's own arrive.expect_tx; bytes arrive from both CTAs' TMA loads
      mbar_init(&empty_bar[s], 1); // multicast tcgen05.commit
    }
    mbar_init(&tmem_full_bar, 1);  // multicast tcgen05.commit
    mbar_init(&tmem_empty_bar, 2 * (blockDim.x / 32 - 4)); // (leader's copy is used) one arrive per epilogue warp of both CTAs
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) { // the whole tensor memory of both CTAs: one 256 x BN fp32 accumulator
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  tc_fence_before();
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (both CTAs): this CTA's half of A and of both W column halves =====
    int it = 0;
    // programmatic dependent launch: the weights do not depend on the preceding kernel
    int pre = 0;
    if (pair < tiles) {
      int const n_blk = pair / m_tiles;
      pre = min(W5_STAGES, num_k);
      for (int i = 0; i < pre; i++) {
        uint32_t const leader_full = w5_mapa(smem_u32(&full_bar[i]), 0);
        if (rank == 0) {
          mbar_expect_tx(&full_bar[i], 2 * W5_STAGE);
        }
        uint32_t const sb = smem_u32(smem + i * W5_STAGE + W5_A_BYTES);
        w5_tma_load(sb, &tmB, i * W5_BK, n_blk * BN + (int)rank * (H1 / 2), leader_full);
        w5_tma_load(sb + B1_BYTES, &tmB2, i * W5_BK, n_blk * BN + H1 + (int)rank * (H2 / 2), leader_full);
      }
    }
    pdl_wait();
    pdl_trigger();
    for (int t = pair; t < tiles; t += P) {
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      for (int kb = 0; kb < num_k; kb++, it++) {
        int const s = it % W5_STAGES;
        uint32_t const ph = (it / W5_STAGES) & 1;
        bool const first_fill = it < pre;
        uint32_t const leader_full = w5_mapa(smem_u32(&full_bar[s]), 0);
        uint32_t const sa = smem_u32(smem + s * W5_STAGE);
        if (!first_fill) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (rank == 0) {
            mbar_expect_tx(&full_bar[s], 2 * W5_STAGE);
          }
        }
        w5_tma_load(sa, &tmA, kb * W5_BK, m_blk * 256 + (int)rank * W5_BM, leader_full);
        if (!first_fill) {
          w5_tma_load(sa + W5_A_BYTES, &tmB, kb * W5_BK, n_blk * BN + (int)rank * (H1 / 2), leader_full);
          w5_tma_load(sa + W5_A_BYTES + B1_BYTES, &tmB2, kb * W5_BK, n_blk * BN + H1 + (int)rank * (H2 / 2), leader_full);
        }
      }
    }
  } else if (warp == 1 && lane == 0) {
    if (rank == 0) {
      // ===== MMA issuer (leader CTA only): UMMA 256 x H1 x 16 and 256 x H2 x 16 per k-step =====
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      uint32_t const idesc0 = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(256 >> 4) << 24);
      uint32_t const idesc1 = idesc0 | ((uint32_t)(H1 >> 3) << 17), idesc2 = idesc0 | ((uint32_t)(H2 >> 3) << 17);
      int it = 0, i = 0;
      for (int t = pair; t < tiles; t += P, i++) {
        mbar_wait(&tmem_empty_bar, (i & 1) ^ 1); // both CTAs' epilogues have drained the accumulator
        tc_fence_after();
        for (int kb = 0; kb < num_k; kb++, it++) {
          int const s = it % W5_STAGES;
          mbar_wait(&full_bar[s], (it / W5_STAGES) & 1);
          tc_fence_after();
          uint32_t const sa = smem_u32(smem + s * W5_STAGE);
          uint64_t const da = umma_desc_sw128(sa);
          uint64_t const db0 = umma_desc_sw128(sa + W5_A_BYTES), db1 = umma_desc_sw128(sa + W5_A_BYTES + B1_BYTES);
#pragma unroll
          for (int k = 0; k < W5_BK / 16; k++) {
            w5_mma(tmem_base, da + (uint64_t)(k * 2), db0 + (uint64_t)(k * 2), idesc1, (kb | k) != 0);
            w5_mma(tmem_base + (uint32_t)H1, da + (uint64_t)(k * 2), db1 + (uint64_t)(k * 2), idesc2, (kb | k) != 0);
          }
          w5_commit(&empty_bar[s]);
        }
        w5_commit(&tmem_full_bar);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): this CTA's 128 x BN half of the accumulator; two warps per 32-lane quarter of tensor
    // memory, one takes the even 64-column chunks, the other the odd ones =====
    int const q = warp & 3, c_first = ((warp - 4) >> 2) * 64, c_step = (int)(blockDim.x / 128 - 1) * 64;
    uint8_t *const stg = stage_out + (warp - 4) * 4096;
    int i = 0;
    for (int t = pair; t < tiles; t += P, i++) {
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      mbar_wait(&tmem_full_bar, i & 1);
      tc_fence_after();
      int const row0 = m_blk * 256 + (int)rank * W5_BM + q * 32, col0 = n_blk * BN;
      uint32_t const tsrc = tmem_base + ((uint32_t)(q * 32) << 16);
      uint32_t ra[32], rb[32];
      tmem_ld32_nowait(tsrc + (uint32_t)c_first, ra); // (BN >= 272: both warps of a quarter have chunks)
      tmem_ld32_nowait(tsrc + (uint32_t)c_first + 32u, rb);
#pragma unroll 1
      for (int c0 = c_first; c0 < BN; c0 += c_step) { // (columns past BN / N are loaded, inside the 512 allocated, never stored)
        tmem_ld_wait32(ra);
        reg_tie32(rb);
        T const *const bp = bias ? bias + col0 + c0 : nullptr;
#pragma unroll
        for (int j = 0; j < 8; j++) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; e++) {
            f[e] = __uint_as_float(j < 4 ? ra[j * 8 + e] : rb[(j - 4) * 8 + e]);
          }
          if (bp && c0 + j * 8 < BN && col0 + c0 + j * 8 < N) {
#pragma unroll
            for (int e = 0; e < 8; e++) {
              f[e] += DT<T>::to_f(bp[j * 8 + e]);
            }
          }
          *reinterpret_cast<uint4 *>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = pack8<T>(f);
        }
        if (c0 + c_step < BN) {
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + c_step), ra);
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + c_step + 32), rb);
        }
        __syncwarp();
        int const j = lane & 7, gcol = col0 + c0 + j * 8;
        bool const col_ok = c0 + j * 8 < BN && gcol < N;
#pragma unroll
        for (int ii = 0; ii < 8; ii++) {
          int const r = ii * 4 + (lane >> 3);
          uint4 const v = *reinterpret_cast<uint4 const *>(stg + r * 128 + ((j ^ (r & 7)) << 4));
          if (col_ok && row0 + r < M) {
            *reinterpret_cast<uint4 *>(out + (size_t)(row0 + r) * N + gcol) = v;
          }
        }
        __syncwarp();
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        uint32_t const remote = w5_mapa(smem_u32(&tmem_empty_bar), 0);
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
    }
  }
  tc_fence_before();
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

template <typename T>
int launch_w5(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN, cudaStream_t s) {
  int const H1 = (BN / 2 + 31) / 32 * 32, H2 = BN - H1; // both multiples of 16 (BN % 16 == 0), 112 <= H2 <= H1 <= 256
  CUtensorMap tmA, tmB, tmB2;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, W5_BM);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, H1 / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB2, dtype, w, N, K, H2 / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  size_t const stage = W5_A_BYTES + (size_t)(BN / 2) * W5_BK * 2;
  int const stages = std::min<int>(W5_MAX_STAGES, (int)(W5_RING_BYTES / stage));
  size_t const smem = (size_t)stages * stage + W5_STAGING_BYTES + 1024; // ring + epilogue staging + alignment slack
  FFB_SET_MAX_SMEM_ONCE(linear_w512_kernel<T>, W5_RING_BYTES + W5_STAGING_BYTES + 1024);
  int const tiles = ((M + 255) / 256) * ((N + BN - 1) / BN);
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * std::min(tiles, kNumSMs / 2));
  static int const epi_sets = env("FFB_W512_EPI") && atoi(env("FFB_W512_EPI")) == 1 ? 1 : 2; // 4 or 8 epilogue warps
  cfg.blockDim = dim3(128 + 128 * epi_sets);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = gemm_pdl_enabled() ? 1 : 0;
  FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, linear_w512_kernel<T>, tmA, tmB, tmB2, (T *)out, (T const *)bias, M, N, K, BN, H1, stages));
  return FFB_OK;
}

} // namespace

// The narrowest tile width (a multiple of 16 in [272, 512]) whose tiles all fit ONE wave of SM pairs; 0 if none does
// (or FFB_W512_BN, any width in range, then possibly several tiles per pair).
int linear_w512_pick_bn(int M, int N) {
  if (char const *e = env("FFB_W512_BN")) {
    int const v = atoi(e);
    if (v >= 272 && v <= 512 && v % 16 == 0) {
      return v;
    }
  }
  int const pairs = kNumSMs / 2, mt = (M + 255) / 256;
  for (int c = 272; c <= 512; c += 16) {
    if (mt * ((N + c - 1) / c) <= pairs) {
      return c;
    }
  }
  return 0;
}

// Selected when 256-wide tiles do not fit one wave of pairs but a wider tile does and fills most of it (M = 512:
// N = 12288 -> BN = 336, 74 tiles, 42.2 against 45.4 us for the stream-K kernel; N = 11008 -> 304; M = 256, N = 32000 ->
// 448, 55.4 against 60.0 us).  With few tiles (o_proj: 32 tiles of 272) the k-split of linear_tc2 is far better.
// FFB_W512=1 forces it wherever a width qualifies, =0 disables it.
int linear_w512_supported(int dtype, int M, int K, int N) {
  if (!((dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && M > 128 && K % W5_BK == 0 && N % 16 == 0)) {
    return 0;
  }
  int const bn = linear_w512_pick_bn(M, N);
  if (bn == 0) {
    return 0;
  }
  if (char const *e = env("FFB_W512")) {
    return e[0] == '1';
  }
  int const pairs = kNumSMs / 2, mt = (M + 255) / 256;
  return mt * ((N + 255) / 256) > pairs && 10 * mt * ((N + bn - 1) / bn) >= 8 * pairs;
}

int launch_linear_w512(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                       cudaStream_t s) {
  int const bn = linear_w512_pick_bn(M, N);
  if (bn == 0) {
    set_error("linear_w512: no tile width fits M=%d N=%d", M, N);
    return FFB_ERR_UNSUPPORTED;
  }
  if (dtype == FFB_DT_HALF) {
    return launch_w5<__half>(in, out, w, bias, dtype, M, K, N, bn, s);
  }
  return launch_w5<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, s);
}

} // namespace ffb
