// linear_tc2.cu — 2-SM (cta_group::2) tcgen05 / TMA projection GEMM for batched prompt / tree tokens (M > 128).
// Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:507-528): out[M][N] = A[M][K] * W[N][K]^T, both operands K-major.
//
// The 1-SM kernel (linear_tc.cu, 128 x 128 tiles) moves (128 + 128) x K x 2 bytes from L2 per 128 x 128 x K x 2
// flops and saturates the L2 -> SM fabric at ~740 TFLOP/s.  Here a cluster of two CTAs (one SM pair) owns a
// 256 x BN output tile: each CTA loads ITS 128 rows of A and ITS BN/2 rows of W per k-block (TMA, 128B swizzle,
// completion signalled on the leader CTA's mbarrier), the leader's elected thread issues ONE
// tcgen05.mma.cta_group::2 (UMMA 256 x BN x 16) that reads both CTAs' shared memory, and each CTA keeps its
// 128 x BN half of the fp32 accumulator in its own tensor memory.  Per-SM operand traffic drops to
// (128 + BN/2) x K x 2 bytes per 128 x BN x K x 2 flops (half of the 1-SM kernel at BN = 256, 2/3 at BN = 128).
// tcgen05.commit (multicast to both CTAs) releases the ring slots and publishes the accumulator; the four
// epilogue warps of each CTA read it with tcgen05.ld, add the bias, convert and store.
// The reference accumulates in DT (cuBLAS compute type = output type); fp32 here.
#include <algorithm>
#include <type_traits>

#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int T2_BM = 128, T2_BK = 64, T2_THREADS = 256;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// shared::cluster address of `local` (a shared::cta address of this CTA) in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t mapa_u32(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
// TMA load into THIS CTA's shared memory whose completion is counted on an mbarrier of the pair's leader CTA
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, CUtensorMap const *tm, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tc_mma_f16_2sm(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// arrive (once all prior MMAs of this thread have completed) on the barrier at the same offset in both CTAs
__device__ __forceinline__ void tc_commit_2sm(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// BN (tile width, a multiple of 16 up to 256) and the ring depth are run-time values: the host picks the width that
// fills the 74 SM pairs in whole waves (e.g. 168 for M = 512, N = 12288: 2 x 74 tiles).
// PERSISTENT: cluster c works on tiles c, c + #clusters, ...; the accumulator is double-buffered in tensor memory
// (2 x 256 columns), so the epilogue of tile i (TMEM -> registers -> global) overlaps the main loop of tile i + 1
// and the launch / barrier / TMEM set-up cost is paid once per SM pair, not once per tile.
constexpr int T2_MAX_STAGES = 10;

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T2_THREADS, 1)
linear_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K, int BN, int STAGES) {
  constexpr uint32_t A_BYTES = T2_BM * T2_BK * 2;
  uint32_t const B_BYTES = (uint32_t)(BN / 2) * T2_BK * 2; // a multiple of 1024 (BN % 16 == 0)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[T2_MAX_STAGES], empty_bar[T2_MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  // 128B-swizzled tiles need 1024-byte alignment; both CTAs compute the same offsets (the MMA addresses the peer's
  // shared memory with the leader's descriptors)
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + STAGES * A_BYTES;

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = cluster_ctarank(); // 0 = leader
  int const num_k = K / T2_BK;
  int const m_tiles = (M + 255) / 256, n_tiles = (N + BN - 1) / BN, tiles = m_tiles * n_tiles;
  int const cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);  // leader's own arrive.expect_tx; bytes arrive from both CTAs' TMA loads
      mbar_init(&empty_bar[s], 1); // multicast tcgen05.commit
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(&tmem_full_bar[b], 1);  // multicast tcgen05.commit
      mbar_init(&tmem_empty_bar[b], 8); // (leader's copy is used) one arrive per epilogue warp of both CTAs
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) { // two accumulator buffers of 256 fp32 columns (BN used) in BOTH CTAs of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  tc_fence_before();
  cluster_sync_all(); // barriers of both CTAs are initialised before any remote arrive / TMA completion
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (both CTAs): this CTA's half of A and of W =====
    int it = 0; // ring item counter, runs across tiles
    for (int t = cluster_id; t < tiles; t += num_clusters) {
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      for (int kb = 0; kb < num_k; kb++, it++) {
        int const s = it % STAGES;
        uint32_t const ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1);
        uint32_t const leader_full = mapa_u32(smem_u32(&full_bar[s]), 0);
        if (rank == 0) {
          mbar_expect_tx(&full_bar[s], 2 * (A_BYTES + B_BYTES));
        }
        tma_load_2d_2sm(smem_u32(sA + s * A_BYTES), &tmA, kb * T2_BK, m_blk * 256 + (int)rank * T2_BM, leader_full);
        tma_load_2d_2sm(smem_u32(sB + s * B_BYTES), &tmB, kb * T2_BK, n_blk * BN + (int)rank * (BN / 2), leader_full);
      }
    }
  } else if (warp == 1 && lane == 0) {
    if (rank == 0) {
      // ===== MMA issuer (leader CTA only) =====
      // cute::UMMA::InstrDescriptor: c_format F32 (1) @4, a/b format @7/@10 (F16 0, BF16 1), K-major, N>>3 @17, M>>4 @24
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      uint32_t const idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
      int it = 0, i = 0;
      for (int t = cluster_id; t < tiles; t += num_clusters, i++) {
        int const b = i & 1;
        mbar_wait(&tmem_empty_bar[b], ((i >> 1) & 1) ^ 1); // both CTAs' epilogues have drained this buffer
        tc_fence_after();
        uint32_t const tmem_d = tmem_base + (uint32_t)b * 256u;
        for (int kb = 0; kb < num_k; kb++, it++) {
          int const s = it % STAGES;
          uint32_t const ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          uint64_t const da = umma_desc_sw128(smem_u32(sA + s * A_BYTES));
          uint64_t const db = umma_desc_sw128(smem_u32(sB + s * B_BYTES));
#pragma unroll
          for (int k = 0; k < T2_BK / 16; k++) {
            // advance 16 elements (32 B) along K inside the swizzle atom: start-address field += 2
            tc_mma_f16_2sm(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          }
          tc_commit_2sm(&empty_bar[s]); // slot reusable in BOTH CTAs once these MMAs have read it
        }
        tc_commit_2sm(&tmem_full_bar[b]);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): this CTA's 128 x BN half of the accumulator, TMEM -> registers -> global =====
    int const q = warp & 3; // this warp may touch TMEM lanes [32q, 32q+32)
    int i = 0;
    for (int t = cluster_id; t < tiles; t += num_clusters, i++) {
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      int const b = i & 1;
      mbar_wait(&tmem_full_bar[b], (i >> 1) & 1);
      tc_fence_after();
      int const row = m_blk * 256 + (int)rank * T2_BM + q * 32 + lane;
      uint32_t const tsrc = tmem_base + (uint32_t)b * 256u + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += 16) {
        uint32_t r[16];
        tmem_ld16(tsrc + (uint32_t)c0, r);
        int const col = n_blk * BN + c0;
        if (row < M && col < N) { // N % 16 == 0: a 16-column chunk is entirely inside or outside the matrix
          T *dst = out + (size_t)row * N + col;
#pragma unroll
          for (int v = 0; v < 2; v++) {
            float f[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
              f[e] = __uint_as_float(r[v * 8 + e]);
              if (bias) {
                f[e] += DT<T>::to_f(bias[col + v * 8 + e]);
              }
            }
            *reinterpret_cast<uint4 *>(dst + v * 8) = pack8<T>(f);
          }
        }
      }
      // hand the buffer back to the leader's MMA issuer
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        uint32_t const remote = mapa_u32(smem_u32(&tmem_empty_bar[b]), 0);
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
    }
  }
  tc_fence_before();
  cluster_sync_all(); // neither CTA frees tensor memory / exits while the peer may still use its smem or TMEM
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

template <typename T>
int launch_one(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN,
               cudaStream_t s) {
  size_t const stage = (size_t)(T2_BM + BN / 2) * T2_BK * 2;
  int const stages = std::min<int>(T2_MAX_STAGES, (int)(((size_t)196 << 10) / stage));
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, T2_BM);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, BN / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  static bool attr_set = false;
  if (!attr_set) {
    FFB_CUDA_TRY(cudaFuncSetAttribute(linear_tc2_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((size_t)197 << 10) + 1024)));
    attr_set = true;
  }
  size_t const smem = (size_t)stages * stage + 1024;
  int const tiles = ((N + BN - 1) / BN) * ((M + 255) / 256);
  dim3 grid(2 * std::min(tiles, kNumSMs / 2));
  linear_tc2_kernel<T><<<grid, T2_THREADS, smem, s>>>(tmA, tmB, (T *)out, (T const *)bias, M, N, K, BN, stages);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace

int linear_tc2_supported(int dtype, int M, int K, int N) {
  if (getenv("FFB_DISABLE_TC2")) {
    return 0;
  }
  return (dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && M > 0 && K % T2_BK == 0 && N % 16 == 0;
}

// w: dense [N][K] DT weights (already dequantised for quantised layers)
int launch_linear_tc2(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                      cudaStream_t s) {
  // Tile width.  The kernel is bound by the L2 -> SM fabric (~6300 B/cycle chip-wide, B300_MICROARCH.md): with all
  // 74 SM pairs busy a tile costs ~(256 + BN) operand rows per k-block, and a pair running alone is only ~12 %
  // faster (measured), so a wave of `a` tiles costs (256 + BN) * max(a / 74, 0.885).  Pick the multiple of 16 that
  // minimises the sum over waves - e.g. 168 for M = 512, N = 12288 (2 x 74 tiles = exactly two full waves).
  int const pairs = kNumSMs / 2;
  long const mt = (M + 255) / 256;
  int bn = 128;
  double best = -1.0;
  for (int c = 64; c <= 256; c += 16) {
    long const tiles = mt * ((N + c - 1) / c);
    long const full = tiles / pairs, rem = tiles % pairs;
    double const cost = (256.0 + c) * ((double)full + (rem ? std::max((double)rem / pairs, 0.885) : 0.0));
    if (best < 0 || cost < best - 1e-9) {
      best = cost;
      bn = c;
    }
  }
  if (char const *e = getenv("FFB_TC2_BN")) {
    int v = atoi(e);
    if (v >= 16 && v <= 256 && v % 16 == 0) {
      bn = v;
    }
  }
  if (dtype == FFB_DT_HALF) {
    return launch_one<__half>(in, out, w, bias, dtype, M, K, N, bn, s);
  }
  return launch_one<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, s);
}

} // namespace ffb
