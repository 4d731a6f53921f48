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
#include <cstdio>
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

// BN (tile width, a multiple of 16 up to 256) and the ring depth are run-time values: the host picks the width that
// fills the 74 SM pairs in whole waves (e.g. 168 for M = 512, N = 12288: 2 x 74 tiles).
// PERSISTENT: cluster c works on tiles c, c + #clusters, ...; the accumulator is double-buffered in tensor memory
// (2 x 256 columns), so the epilogue of tile i (TMEM -> registers -> global) overlaps the main loop of tile i + 1
// and the launch / barrier / TMEM set-up cost is paid once per SM pair, not once per tile.
constexpr int T2_MAX_STAGES = 10;

// split-K work unit u = (tile u / S, k-slice u % S): the slices of a tile, which wait for each other in the epilogue,
// run on neighbouring clusters (scheduled together)
__device__ __forceinline__ int unit_slice(int u, int S) { return S == 1 ? 0 : u % S; }

#ifdef FFB_TC2_TRACE_BUILD
// nvcc -DFFB_TC2_TRACE_BUILD + env FFB_TC2_TRACE=1: the leader CTA of cluster 0 stamps %globaltimer at its phase
// boundaries for 8 consecutive launches (row = launch % 8); dumped by the host after every 8th launch
__device__ long long g_tc2_trace[8][16];
__device__ int g_tc2_launch;
#define T2_TR(ev)                                                                  \
  do {                                                                             \
    if (blockIdx.x == 0) {                                                         \
      long long t_;                                                                \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                       \
      g_tc2_trace[*(volatile int *)&g_tc2_launch & 7][ev] = t_;                    \
    }                                                                              \
  } while (0)
#else
#define T2_TR(ev) ((void)0)
#endif

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(T2_THREADS, 1)
linear_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K, int BN, int STAGES,
                  int S, float4 *__restrict__ ws, int *__restrict__ cnt) {
  constexpr uint32_t A_BYTES = T2_BM * T2_BK * 2;
  uint32_t const B_BYTES = (uint32_t)(BN / 2) * T2_BK * 2; // a multiple of 1024 (BN % 16 == 0)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[T2_MAX_STAGES], empty_bar[T2_MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t part_bar[4][4]; // split-K: partial of epilogue warp q, 64-column chunk c has landed
  __shared__ uint32_t tmem_base_smem;
  // 128B-swizzled tiles need 1024-byte alignment; both CTAs compute the same offsets (the MMA addresses the peer's
  // shared memory with the leader's descriptors)
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + STAGES * A_BYTES;
  uint8_t *stage_out = sB + STAGES * B_BYTES; // 4 epilogue warps x 4 KB (a multiple of 1024 from smem)

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = cluster_ctarank(); // 0 = leader
  int const num_k = K / T2_BK;
  int const m_tiles = (M + 255) / 256, n_tiles = (N + BN - 1) / BN, tiles = m_tiles * n_tiles;
  int const cluster_id = blockIdx.x >> 1, num_clusters = gridDim.x >> 1;
  // split-K: work unit u = (tile u / S, k-slice u % S); slice ks covers k-blocks [ks*num_k/S, (ks+1)*num_k/S)
  int const units = tiles * S;

  if (warp == 0 && lane == 0) {
    T2_TR(0);
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
    for (int i = 0; i < 16; i++) {
      mbar_init(&part_bar[i >> 2][i & 3], 1);
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
    T2_TR(1);
    // Launched as a programmatic dependent of the preceding kernel: the weights do not depend on it, so the W halves of
    // the first ring fill (k-blocks of this cluster's first unit) are requested at once; the activations wait.
    int pre = 0;
    if (cluster_id < units) {
      int const t = cluster_id / S, ks = unit_slice(cluster_id, S);
      int const n_blk = t / m_tiles, kb0 = ks * num_k / S;
      pre = min(STAGES, (ks + 1) * num_k / S - kb0);
      for (int i = 0; i < pre; i++) {
        uint32_t const leader_full = mapa_u32(smem_u32(&full_bar[i]), 0);
        if (rank == 0) {
          mbar_expect_tx(&full_bar[i], 2 * (A_BYTES + B_BYTES));
        }
        tma_load_2d_2sm(smem_u32(sB + i * B_BYTES), &tmB, (kb0 + i) * T2_BK, n_blk * BN + (int)rank * (BN / 2), leader_full);
      }
    }
    pdl_wait();
    pdl_trigger();
    for (int u = cluster_id; u < units; u += num_clusters) {
      int const t = u / S, ks = unit_slice(u, S);
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      int const kb1 = (ks + 1) * num_k / S;
      for (int kb = ks * num_k / S; kb < kb1; kb++, it++) {
        int const s = it % STAGES;
        uint32_t const ph = (it / STAGES) & 1;
        bool const first_fill = it < pre;
        uint32_t const leader_full = mapa_u32(smem_u32(&full_bar[s]), 0);
        if (!first_fill) {
          mbar_wait(&empty_bar[s], ph ^ 1);
          if (rank == 0) {
            mbar_expect_tx(&full_bar[s], 2 * (A_BYTES + B_BYTES));
          }
        }
        tma_load_2d_2sm(smem_u32(sA + s * A_BYTES), &tmA, kb * T2_BK, m_blk * 256 + (int)rank * T2_BM, leader_full);
        if (!first_fill) {
          tma_load_2d_2sm(smem_u32(sB + s * B_BYTES), &tmB, kb * T2_BK, n_blk * BN + (int)rank * (BN / 2), leader_full);
        }
      }
    }
    T2_TR(2);
  } else if (warp == 1 && lane == 0) {
    if (rank == 0) {
      // ===== MMA issuer (leader CTA only) =====
      // cute::UMMA::InstrDescriptor: c_format F32 (1) @4, a/b format @7/@10 (F16 0, BF16 1), K-major, N>>3 @17, M>>4 @24
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      uint32_t const idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
      int it = 0, i = 0;
      for (int u = cluster_id; u < units; u += num_clusters, i++) {
        int const b = i & 1;
        mbar_wait(&tmem_empty_bar[b], ((i >> 1) & 1) ^ 1); // both CTAs' epilogues have drained this buffer
        tc_fence_after();
        uint32_t const tmem_d = tmem_base + (uint32_t)b * 256u;
        int const ks = unit_slice(u, S), kb0 = ks * num_k / S, kb1 = (ks + 1) * num_k / S;
        for (int kb = kb0; kb < kb1; kb++, it++) {
          int const s = it % STAGES;
          uint32_t const ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          if (it == 0) {
            T2_TR(3);
          }
          tc_fence_after();
          uint64_t const da = umma_desc_sw128(smem_u32(sA + s * A_BYTES));
          uint64_t const db = umma_desc_sw128(smem_u32(sB + s * B_BYTES));
#pragma unroll
          for (int k = 0; k < T2_BK / 16; k++) {
            // advance 16 elements (32 B) along K inside the swizzle atom: start-address field += 2
            tc_mma_f16_2sm(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, ((kb - kb0) | k) != 0);
          }
          tc_commit_2sm(&empty_bar[s]); // slot reusable in BOTH CTAs once these MMAs have read it
        }
        tc_commit_2sm(&tmem_full_bar[b]);
      }
      T2_TR(4);
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): this CTA's 128 x BN half of the accumulator, TMEM -> registers -> global =====
    int const q = warp & 3; // this warp may touch TMEM lanes [32q, 32q+32)
    int i = 0;
    for (int u = cluster_id; u < units; u += num_clusters, i++) {
      int const t = u / S;
      int const m_blk = t % m_tiles, n_blk = t / m_tiles;
      int const b = i & 1;
      mbar_wait(&tmem_full_bar[b], (i >> 1) & 1);
      if (warp == 4 && lane == 0) {
        T2_TR(5);
      }
      tc_fence_after();
      int const row = m_blk * 256 + (int)rank * T2_BM + q * 32 + lane;
      uint32_t const tsrc = tmem_base + (uint32_t)b * 256u + ((uint32_t)(q * 32) << 16);
      // ----- split-K (S == 2): the two k-slices of a tile run on neighbouring SM pairs and EXCHANGE halves: slice 0
      // finishes the low 64-column chunks of the tile, slice 1 the high ones.  Each warp first parks the chunks it
      // does not finish as fp32 in the L2 workspace (per 16 columns: 4 x 32 lanes x float4, so that its stores and
      // the partner's bulk copies are contiguous) and raises a flag; then it waits for the partner warp (same rows),
      // bulk-copies the partner's partial of ITS chunks into the operand ring - idle by now: every MMA of this
      // CTA's only unit has completed - one mbarrier per chunk, and adds it to the accumulator below.  A two-term
      // fp32 sum does not depend on which side adds, so the result is deterministic. -----
      int const nchunk = (BN + 63) >> 6;
      int c_begin = 0, c_end = nchunk;
      uint8_t *const part = smem + (size_t)q * (BN >> 4) * 2048;
      if (S == 2) {
        int const ks = unit_slice(u, S), wid = (int)rank * 4 + q, half = nchunk >> 1;
        c_begin = ks ? half : 0;
        c_end = ks ? nchunk : half;
        int const o_begin = ks ? 0 : half, o_end = ks ? half : nchunk;
        float4 *const my = ws + ((size_t)((t * 2 + ks) * 8 + wid) * (BN >> 4)) * 128 + lane;
        uint32_t ra[32], rb[32];
        if (o_begin < o_end) {
          tmem_ld32_nowait(tsrc + (uint32_t)(o_begin * 64), ra);
          tmem_ld32_nowait(tsrc + (uint32_t)(o_begin * 64 + 32), rb);
        }
#pragma unroll 1
        for (int c = o_begin; c < o_end; c++) {
          int const c0 = c * 64;
          tmem_ld_wait32(ra);
          reg_tie32(rb);
#pragma unroll
          for (int j = 0; j < 16; j++) {
            if (c0 + j * 4 < BN) {
              uint32_t const *r = j < 8 ? &ra[j * 4] : &rb[(j - 8) * 4];
              __stcg(my + ((c0 >> 2) + j) * 32, make_float4(__uint_as_float(r[0]), __uint_as_float(r[1]),
                                                            __uint_as_float(r[2]), __uint_as_float(r[3])));
            }
          }
          if (c + 1 < o_end) {
            tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 64), ra);
            tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 96), rb);
          }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) {
          atomicExch(&cnt[(t * 2 + ks) * 8 + wid], 1);
          int *const theirs = cnt + (t * 2 + (ks ^ 1)) * 8 + wid;
          int seen;
          do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(theirs) : "memory");
          } while (seen == 0);
          *theirs = 0; // ready for the next launch on this stream (the partner only writes its flag once)
          asm volatile("fence.proxy.async;" ::: "memory"); // the partials were written through the generic proxy
          float4 const *const src = ws + ((size_t)((t * 2 + (ks ^ 1)) * 8 + wid) * (BN >> 4)) * 128;
          for (int c = c_begin; c < c_end; c++) {
            uint32_t const bytes = (uint32_t)(min(64, BN - c * 64) >> 4) * 2048u;
            mbar_expect_tx(&part_bar[q][c], bytes);
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(smem_u32(part + (size_t)c * 8192)), "l"(src + (size_t)c * 512), "r"(bytes),
                           "r"(smem_u32(&part_bar[q][c])) : "memory");
          }
        }
        __syncwarp();
      }
      // 64-column chunks: TMEM -> registers (+ bias, convert) -> this warp's 32 x 128 B staging tile (16-byte units
      // XOR-swizzled by row, conflict-free both ways) -> global with 8 lanes per row, so that every warp store
      // covers 4 full 128-byte lines (a thread-per-row store touches 32 lines per instruction and made the
      // epilogue LSU-bound: 2.4 us per 256-column tile).  The tcgen05.ld of chunk c + 1 is in flight during the
      // read-back of chunk c.  Columns past BN / N are loaded (inside the 256-column buffer) but never stored.
      int const col0 = n_blk * BN, row0 = row - lane;
      uint8_t *const stg = stage_out + (warp - 4) * 4096;
      uint32_t ra[32], rb[32];
      if (c_begin < c_end) {
        tmem_ld32_nowait(tsrc + (uint32_t)(c_begin * 64), ra);
        tmem_ld32_nowait(tsrc + (uint32_t)(c_begin * 64 + 32), rb);
      }
#pragma unroll 1
      for (int c0 = c_begin * 64; c0 < c_end * 64; c0 += 64) {
        tmem_ld_wait32(ra);
        reg_tie32(rb);
        if (S == 2) {
          mbar_wait(&part_bar[q][c0 >> 6], 0);
          float4 const *p4 = reinterpret_cast<float4 const *>(part + (size_t)(c0 >> 6) * 8192) + lane;
#pragma unroll
          for (int j = 0; j < 16; j++) {
            if (c0 + j * 4 < BN) {
              float4 const a = p4[j * 32];
              uint32_t *r = j < 8 ? &ra[j * 4] : &rb[(j - 8) * 4];
              r[0] = __float_as_uint(__uint_as_float(r[0]) + a.x);
              r[1] = __float_as_uint(__uint_as_float(r[1]) + a.y);
              r[2] = __float_as_uint(__uint_as_float(r[2]) + a.z);
              r[3] = __float_as_uint(__uint_as_float(r[3]) + a.w);
            }
          }
        }
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
        if (c0 + 64 < c_end * 64) {
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 64), ra);
          tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 96), rb);
        }
        __syncwarp();
        int const j = lane & 7, gcol = col0 + c0 + j * 8;
        bool const col_ok = gcol < N && c0 + j * 8 < BN;
#pragma unroll
        for (int i = 0; i < 8; i++) {
          int const r = i * 4 + (lane >> 3);
          uint4 const v = *reinterpret_cast<uint4 const *>(stg + r * 128 + ((j ^ (r & 7)) << 4));
          if (col_ok && row0 + r < M) {
            *reinterpret_cast<uint4 *>(out + (size_t)(row0 + r) * N + gcol) = v;
          }
        }
        __syncwarp();
      }
      if (warp == 4 && lane == 0) {
        T2_TR(6);
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
  // neither CTA frees tensor memory / exits while the peer may still use its smem or TMEM; execution ordering only
  // (a .release arrive would wait for this thread's global stores to drain: 1 us)
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
#ifdef FFB_TC2_TRACE_BUILD
  if (threadIdx.x == 0) {
    T2_TR(7);
    if (blockIdx.x == 0) {
      __threadfence();
      g_tc2_launch = g_tc2_launch + 1;
    }
  }
#endif
}

template <typename T>
int launch_one(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN,
               int S, void *ws, int *cnt, cudaStream_t s) {
  size_t const stage = (size_t)(T2_BM + BN / 2) * T2_BK * 2;
  int const stages = std::min<int>(T2_MAX_STAGES, (int)(((size_t)208 << 10) / stage));
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, T2_BM);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, BN / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  FFB_SET_MAX_SMEM_ONCE(linear_tc2_kernel<T>, ((size_t)225 << 10) + 1024);
  size_t const smem = (size_t)stages * stage + 16384 + 1024; // ring + epilogue staging + alignment slack
  int const tiles = ((N + BN - 1) / BN) * ((M + 255) / 256);
  dim3 grid(2 * std::min(tiles * S, kNumSMs / 2));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = dim3(T2_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = gemm_pdl_enabled() ? 1 : 0;
  FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, linear_tc2_kernel<T>, tmA, tmB, (T *)out, (T const *)bias, M, N, K, BN, stages, S,
                                  (float4 *)ws, cnt));
#ifdef FFB_TC2_TRACE_BUILD
  static int calls = 0;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(s, &cap);
  if (ffb::env("FFB_TC2_TRACE") && cap == cudaStreamCaptureStatusNone && ++calls % 8 == 0) {
    long long host[8][16];
    cudaStreamSynchronize(s);
    cudaMemcpyFromSymbol(host, g_tc2_trace, sizeof(host));
    for (int r = 0; r < 8; r++) {
      fprintf(stderr, "TC2TRACE M %d K %d N %d BN %d S %d:", M, K, N, BN, S);
      for (int e = 0; e < 8; e++) {
        fprintf(stderr, " %lld", host[r][e] - host[0][0]);
      }
      fprintf(stderr, "\n");
    }
  }
#endif
  return FFB_OK;
}

} // namespace

int linear_tc2_supported(int dtype, int M, int K, int N) {
  if (ffb::env("FFB_DISABLE_TC2")) {
    return 0;
  }
  return (dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) && M > 0 && K % T2_BK == 0 && N % 16 == 0;
}

// split-K workspace: one 256 x 256 fp32 partial per SM pair (S > 1 only when all units fit one wave) + one arrival
// counter per (tile, epilogue warp)
size_t linear_tc2_ws_bytes() { return (size_t)(kNumSMs / 2) * 256 * 256 * 4; }
size_t linear_tc2_cnt_bytes() { return (size_t)(kNumSMs / 2) * 8 * sizeof(int); }

// w: dense [N][K] DT weights (already dequantised for quantised layers); ws / cnt: split-K workspace (or null)
int launch_linear_tc2(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                      void *ws, int *cnt, cudaStream_t s) {
  // Tile width and k-split.  The kernel is bound by the L2 -> SM fabric (~6300 B/cycle chip-wide,
  // B300_MICROARCH.md): with all 74 SM pairs busy a tile costs ~(256 + BN) operand rows per k-block, and a pair
  // running alone is only ~12 % faster (measured), so a wave of `a` units costs (256 + BN) / S * max(a / 74, 0.885).
  // Pick the (multiple of 16, S) that minimises the sum over waves - e.g. BN = 168 for M = 512, N = 12288 (2 x 74
  // tiles = exactly two full waves), BN = 256 with S = 2 for M = 512, N = 4096 (64 units in one wave instead of
  // 32 tiles on 74 pairs).  S > 1 pays a merge (fp32 partials through L2) and must fit one wave.
  int const pairs = kNumSMs / 2, num_k = K / T2_BK;
  long const mt = (M + 255) / 256;
  int bn = 128, split = 1;
  double best = -1.0;
  static double const merge_cost = ffb::env("FFB_TC2_MERGE_COST") ? atof(ffb::env("FFB_TC2_MERGE_COST")) : 64.0;
  for (int S = 1; S <= (ws && cnt ? 2 : 1); S++) {
    if (S > 1 && num_k / S < 16) { // (K = 1024 split in two measured 12.8 against 10.7 us unsplit at M = 512, N = 4096)
      break;
    }
    for (int c = 64; c <= 256; c += 16) {
      long const tiles = mt * ((N + c - 1) / c);
      long const units = tiles * S;
      if (S > 1 && units > pairs) {
        continue;
      }
      long const full = units / pairs, rem = units % pairs;
      double const cost = (256.0 + c) / S * ((double)full + (rem ? std::max((double)rem / pairs, 0.885) : 0.0)) +
                          (S > 1 ? merge_cost : 0.0);
      if (best < 0 || cost < best - 1e-9) {
        best = cost;
        bn = c;
        split = S;
      }
    }
  }
  if (char const *e = ffb::env("FFB_TC2_BN")) {
    int v = atoi(e);
    if (v >= 16 && v <= 256 && v % 16 == 0) {
      bn = v;
    }
  }
  if (char const *e = ffb::env("FFB_TC2_SPLIT")) {
    int v = atoi(e);
    if (v >= 1 && v <= 2 && (v == 1 || (ws && cnt))) { // S - 1 partials of 512 x BN bytes live in the idle ring
      split = v;
    }
  }
  if (split > 1 && (mt * ((N + bn - 1) / bn) * split > pairs || num_k < split)) {
    split = 1; // (env overrides) the workspace holds one wave of units
  }
  if (dtype == FFB_DT_HALF) {
    return launch_one<__half>(in, out, w, bias, dtype, M, K, N, bn, split, ws, cnt, s);
  }
  return launch_one<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, split, ws, cnt, s);
}

} // namespace ffb
