// linear_tc4.cu — the SM-pair (cta_group::2) tcgen05 / TMA projection GEMM of linear_tc2.cu run in CLUSTERS OF TWO
// PAIRS that share their A operand by TMA multicast.  Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:507-528): out[M][N] = A[M][K] * W[N][K]^T.
//
// Why: with few M tiles (512 tree tokens = two 256-row tiles) every SM pair of linear_tc2 re-reads the same A rows for
// its own slice of W, and the kernel runs at the chip's L2 -> SM fabric limit, not at the tensor pipe: M = 512,
// N = 12288, K = 4096 moves 310 MB of A + 200 MB of W per call = 45 us at ~11 TB/s, which is what it takes
// (profiles/README.md).  Here the four CTAs of a cluster form two pairs that work on the SAME 256 rows of A and on
// neighbouring BN-wide column blocks of W: each CTA fetches half of its 128-row A block per k-block and multicasts it
// to the CTA on the same side of the other pair, so a cluster pulls each A byte from L2 once (A traffic / 2); W is loaded
// as before.  The two pairs run their k-loops in lockstep through the ring's empty barriers (a slot is refilled only
// when BOTH pairs' MMAs have read it: the multicast writes into both).  Everything else - UMMA 256 x BN x 16 issued
// by each pair's leader, accumulators double-buffered in tensor memory, staged coalesced epilogue - is linear_tc2's.
#include <algorithm>
#include <cstdio>
#include <type_traits>

#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int T2_BM = 128, T2_BK = 64, T2_THREADS = 256;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu; // shared::cluster address -> the even CTA of the pair (cute: Sm100MmaPeerBitMask)

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
// the same, multicast: the box lands at `dst`'s offset in every CTA of `mask`, and the bytes are counted on the barrier
// at `bar`'s offset in the even CTA of each destination's pair (peer bit cleared, as cute's SM100_TMA_2SM_LOAD_MULTICAST)
__device__ __forceinline__ void tma_load_2d_2sm_mc(uint32_t dst, CUtensorMap const *tm, int c0, int c1, uint32_t bar, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
      ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1), "h"(mask) : "memory");
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
__device__ __forceinline__ void tc_commit_2sm(uint64_t *bar, uint16_t mask) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"(mask) : "memory");
}

// 16 fp32 accumulator columns (+ bias) -> 16 DT values, two 16-byte stores
template <typename T, int OFF = 0, int NR>
__device__ __forceinline__ void emit16(uint32_t const (&r)[NR], T *dst, T const *bias) {
#pragma unroll
  for (int v = 0; v < 2; v++) {
    float f[8];
#pragma unroll
    for (int e = 0; e < 8; e++) {
      f[e] = __uint_as_float(r[OFF + v * 8 + e]);
      if (bias) {
        f[e] += DT<T>::to_f(bias[v * 8 + e]);
      }
    }
    *reinterpret_cast<uint4 *>(dst + v * 8) = pack8<T>(f);
  }
}
// N % 16 == 0: each 16-column half of a 32-column group is entirely inside or outside the matrix
template <typename T>
__device__ __forceinline__ void emit32(uint32_t const (&r)[32], T *orow, T const *bias, int col, bool row_ok, int N) {
  if (row_ok && col < N) {
    emit16<T, 0>(r, orow + col, bias ? bias + col : nullptr);
    if (col + 16 < N) {
      emit16<T, 16>(r, orow + col + 16, bias ? bias + col + 16 : nullptr);
    }
  }
}

// BN (tile width, a multiple of 16 up to 256) and the ring depth are run-time values.  PERSISTENT: cluster c works on
// double tiles c, c + #clusters, ...; double tile u = (M block u % m_tiles, column blocks 2 (u / m_tiles) + pair).
constexpr int T2_MAX_STAGES = 10;

template <typename T>
__global__ void __cluster_dims__(4, 1, 1) __launch_bounds__(T2_THREADS, 1)
linear_tc4_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K, int BN, int STAGES) {
  constexpr uint32_t A_BYTES = T2_BM * T2_BK * 2, A_HALF = A_BYTES / 2;
  uint32_t const B_BYTES = (uint32_t)(BN / 2) * T2_BK * 2; // a multiple of 1024 (BN % 16 == 0)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[T2_MAX_STAGES], empty_bar[T2_MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  // 128B-swizzled tiles need 1024-byte alignment; all four CTAs compute the same offsets (a pair's MMA addresses the
  // peer's shared memory with the leader's descriptors, the multicast writes to the same offset in the other pair)
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + STAGES * A_BYTES;
  uint8_t *stage_out = sB + STAGES * B_BYTES; // 4 epilogue warps x 4 KB

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = cluster_ctarank();           // 0..3
  uint32_t const pair = rank >> 1, side = rank & 1;  // side 0 = the pair's leader (MMA issuer)
  uint32_t const leader = rank & ~1u;
  int const num_k = K / T2_BK;
  int const m_tiles = (M + 255) / 256, n_tiles = (N + BN - 1) / BN; // n_tiles is even (host)
  int const units = m_tiles * (n_tiles / 2);
  int const cluster_id = blockIdx.x >> 2, num_clusters = gridDim.x >> 2;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);  // (pair leaders) own arrive.expect_tx; bytes arrive from the TMA loads of all four CTAs
      mbar_init(&empty_bar[s], 2); // one multicast tcgen05.commit per pair
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(&tmem_full_bar[b], 1);  // the pair's multicast tcgen05.commit
      mbar_init(&tmem_empty_bar[b], 8); // (leaders' copies are used) one arrive per epilogue warp of the pair
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) { // two accumulator buffers of 256 fp32 columns (BN used) in BOTH CTAs of the pair
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
  }
  tc_fence_before();
  cluster_sync_all(); // barriers of all four CTAs are initialised before any remote arrive / TMA completion
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (all four CTAs) =====
    // A: this CTA's 128 rows of the cluster's 256-row block are needed by it and by the same-side CTA of the other pair;
    // each of the two fetches one 64-row half per k-block and multicasts it to both.  W: this CTA's BN/2 rows of its
    // pair's column block, as in linear_tc2.
    uint16_t const a_mask = (uint16_t)((1u << side) | (1u << (2 + side)));
    int it = 0; // ring item counter, runs across tiles
    for (int u = cluster_id; u < units; u += num_clusters) {
      int const m_blk = u % m_tiles, n_blk = 2 * (u / m_tiles) + (int)pair;
      for (int kb = 0; kb < num_k; kb++, it++) {
        int const s = it % STAGES;
        uint32_t const ph = (it / STAGES) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1); // both pairs have read the slot (it is written from both)
        uint32_t const pair_full = smem_u32(&full_bar[s]) & kPeerBitMask; // the even CTA of (each destination's) pair
        if (side == 0) {
          mbar_expect_tx(&full_bar[s], 2 * (A_BYTES + B_BYTES));
        }
        tma_load_2d_2sm_mc(smem_u32(sA + s * A_BYTES + pair * A_HALF), &tmA, kb * T2_BK,
                           m_blk * 256 + (int)side * T2_BM + (int)pair * (T2_BM / 2), pair_full, a_mask);
        tma_load_2d_2sm(smem_u32(sB + s * B_BYTES), &tmB, kb * T2_BK, n_blk * BN + (int)side * (BN / 2),
                        mapa_u32(smem_u32(&full_bar[s]), leader));
      }
    }
  } else if (warp == 1 && lane == 0) {
    if (side == 0) {
      // ===== MMA issuer (each pair's leader) =====
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      uint32_t const idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
      uint16_t const pair_mask = (uint16_t)(3u << (2 * pair));
      int it = 0, i = 0;
      for (int u = cluster_id; u < units; u += num_clusters, i++) {
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
            tc_mma_f16_2sm(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          }
          tc_commit_2sm(&empty_bar[s], (uint16_t)0xF); // slot reusable in ALL FOUR CTAs once both pairs have committed
        }
        tc_commit_2sm(&tmem_full_bar[b], pair_mask);
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (all four CTAs): this CTA's 128 x BN half of its pair's accumulator, TMEM -> registers -> global =====
    int const q = warp & 3; // this warp may touch TMEM lanes [32q, 32q+32)
    int i = 0;
    for (int u = cluster_id; u < units; u += num_clusters, i++) {
      int const m_blk = u % m_tiles, n_blk = 2 * (u / m_tiles) + (int)pair;
      int const b = i & 1;
      mbar_wait(&tmem_full_bar[b], (i >> 1) & 1);
      tc_fence_after();
      int const row = m_blk * 256 + (int)side * T2_BM + q * 32 + lane;
      uint32_t const tsrc = tmem_base + (uint32_t)b * 256u + ((uint32_t)(q * 32) << 16);
      // 64-column chunks: TMEM -> registers (+ bias, convert) -> this warp's 32 x 128 B staging tile (16-byte units
      // XOR-swizzled by row, conflict-free both ways) -> global with 8 lanes per row, so that every warp store covers
      // 4 full 128-byte lines.  The tcgen05.ld of chunk c + 1 is in flight during the read-back of chunk c.  Columns
      // past BN / N are loaded (inside the 256-column buffer) but never stored.
      int const nchunk = (BN + 63) >> 6;
      int const col0 = n_blk * BN, row0 = row - lane;
      uint8_t *const stg = stage_out + (warp - 4) * 4096;
      uint32_t ra[32], rb[32];
      tmem_ld32_nowait(tsrc, ra);
      tmem_ld32_nowait(tsrc + 32u, rb);
#pragma unroll 1
      for (int c0 = 0; c0 < nchunk * 64; c0 += 64) {
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
          if (bp && col0 + c0 + j * 8 < N && c0 + j * 8 < BN) {
#pragma unroll
            for (int e = 0; e < 8; e++) {
              f[e] += DT<T>::to_f(bp[j * 8 + e]);
            }
          }
          *reinterpret_cast<uint4 *>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = pack8<T>(f);
        }
        if (c0 + 64 < nchunk * 64) {
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 64), ra);
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 96), rb);
        }
        __syncwarp();
        int const j = lane & 7, gcol = col0 + c0 + j * 8;
        bool const col_ok = gcol < N && c0 + j * 8 < BN;
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
      // hand the buffer back to the pair leader's MMA issuer
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        uint32_t const remote = mapa_u32(smem_u32(&tmem_empty_bar[b]), leader);
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
    }
  }
  tc_fence_before();
  // no CTA frees tensor memory / exits while another may still use its smem or TMEM; execution ordering only
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

template <typename T>
int launch_tc4(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN, cudaStream_t s) {
  size_t const stage = (size_t)(T2_BM + BN / 2) * T2_BK * 2;
  int const stages = std::min<int>(T2_MAX_STAGES, (int)(((size_t)208 << 10) / stage));
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, T2_BM / 2); // 64-row boxes: half of a CTA's A block per load
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, BN / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  FFB_SET_MAX_SMEM_ONCE(linear_tc4_kernel<T>, ((size_t)225 << 10) + 1024);
  size_t const smem = (size_t)stages * stage + 16384 + 1024; // ring + epilogue staging + alignment slack
  int const units = ((M + 255) / 256) * (((N + BN - 1) / BN) / 2);
  // how many 4-CTA clusters the device holds at once: GPCs whose SM count is not a multiple of 4 lose some (33 of 37 on
  // a 148-SM B200); clusters beyond that would start only when the first ones exit
  static int max_clusters[2] = {0, 0};
  int &mc = max_clusters[dtype == FFB_DT_HALF ? 0 : 1];
  if (mc == 0) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(kNumSMs / 4 * 4);
    cfg.blockDim = dim3(T2_THREADS);
    cfg.dynamicSmemBytes = ((size_t)209 << 10) + 16384 + 1024;
    int nc = 0;
    if (cudaOccupancyMaxActiveClusters(&nc, linear_tc4_kernel<T>, &cfg) != cudaSuccess || nc <= 0) {
      (void)cudaGetLastError();
      nc = kNumSMs / 4 - 5;
    }
    mc = std::min(nc, kNumSMs / 4);
  }
  dim3 grid(4 * std::min(units, mc));
  if (ffb::env("FFB_TC4_DEBUG")) {
    fprintf(stderr, "[ffb] linear_tc4: M=%d N=%d K=%d BN=%d stages=%d units=%d grid=%u max_active_clusters=%d\n", M, N, K, BN, stages,
            units, grid.x, mc);
  }
  linear_tc4_kernel<T><<<grid, T2_THREADS, smem, s>>>(tmA, tmB, (T *)out, (T const *)bias, M, N, K, BN, stages);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace

// Tile width for the clustered kernel: a multiple of 16 with an EVEN number of column blocks, chosen to fill the 37
// clusters in whole waves (cost of a wave ~ max(tensor time ~ BN, fabric time ~ 128 + BN) per k-block; under-filled
// waves are charged as in linear_tc2).  Returns 0 when no width qualifies (the caller keeps linear_tc2).
int linear_tc4_pick_bn(int M, int N) {
  int const clusters = 33; // resident 4-CTA clusters on a 148-SM B200 (launch_tc4 measures it)
  long const mt = (M + 255) / 256;
  int bn = 0;
  double best = -1.0;
  for (int c = 64; c <= 256; c += 16) {
    long const nt = (N + c - 1) / c;
    if (nt % 2 != 0) {
      continue;
    }
    long const units = mt * (nt / 2);
    long const full = units / clusters, rem = units % clusters;
    double const cost = (128.0 + c) * ((double)full + (rem ? std::max((double)rem / clusters, 0.885) : 0.0));
    if (best < 0 || cost < best - 1e-9) {
      best = cost;
      bn = c;
    }
  }
  return bn;
}

int linear_tc4_supported(int dtype, int M, int K, int N) {
  static bool const on = ffb::env("FFB_TC4") != nullptr && atoi(ffb::env("FFB_TC4")) != 0;
  return on && (dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && M > 128 && K % T2_BK == 0 && N % 16 == 0 &&
         linear_tc4_pick_bn(M, N) != 0;
}

int launch_linear_tc4(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N, cudaStream_t s) {
  int bn = linear_tc4_pick_bn(M, N);
  if (char const *e = ffb::env("FFB_TC4_BN")) {
    int const v = atoi(e);
    if (v >= 16 && v <= 256 && v % 16 == 0 && ((N + v - 1) / v) % 2 == 0) {
      bn = v;
    }
  }
  if (dtype == FFB_DT_HALF) {
    return launch_tc4<__half>(in, out, w, bias, dtype, M, K, N, bn, s);
  }
  return launch_tc4<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, s);
}

} // namespace ffb
