// linear_sk.cu — stream-K schedule for the SM-pair (cta_group::2) tcgen05 / TMA projection GEMM of linear_tc2.cu.
// Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel (src/ops/kernels/linear_kernels.cu:507-528) for
// batched prompt / tree tokens: out[M][N] = A[M][K] * W[N][K]^T, both operands K-major, fp32 accumulation.
//
// Why: a 256 x BN x 16 tcgen05.mma costs the same issue slot whatever BN (profiles/README.md), so only 256-wide tiles
// run at the tensor rate - but M = 512, N = 12288 is 96 such tiles for 74 SM pairs: two waves, the second 30 % full
// (51 us; the work is 31 us).  Here the (tile, k-block) space is one flat range of tiles * K/64 k-blocks, tile-major,
// cut into P equal contiguous pieces, one per SM pair.  A pair's piece is: the tail of a tile (k-blocks [k0, K/64)),
// whole tiles, the head of a tile ([0, k1)).  Every pair runs the same number of k-blocks (+-1); the operand ring and
// the double-buffered accumulator run straight through the segment boundaries.
//
// A tile cut by a piece boundary is finished by the pair that holds its HEAD - the last segment of that pair's range -
// from the fp32 partials of the pairs that hold the rest - the FIRST segment of their ranges, so the partials are
// long there when the head is done and no pair ever waits for a pair that waits: each epilogue warp of a writer parks
// its 32 rows x BN columns in an L2 workspace slot (one slot per pair) and raises a flag; the owner's warp of the same
// rows bulk-copies the slot into the idle operand ring and adds it, piece after piece in pair order (a fixed order:
// the result does not depend on timing).  At most ceil((K/64) / piece length) + 1 pieces per tile.
#include <algorithm>
#include <cstdio>
#include <type_traits>

#include "tc_common.cuh"

namespace ffb {
using namespace tc;
namespace {

constexpr int SK_BM = 128, SK_BK = 64, SK_THREADS = 256, SK_MAX_STAGES = 10;

__device__ __forceinline__ uint32_t sk_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ uint32_t sk_mapa(uint32_t local, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(local), "r"(rank));
  return r;
}
// TMA load into THIS CTA's shared memory whose completion is counted on an mbarrier of the pair's leader CTA
__device__ __forceinline__ void sk_tma_load(uint32_t dst, CUtensorMap const *tm, int c0, int c1, uint32_t leader_bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(tm), "r"(leader_bar), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void sk_mma(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
      "}" ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void sk_commit(uint64_t *bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
               ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}

// The piece of pair p: k-blocks [p * total / P, (p + 1) * total / P) of the flat (tile-major) space.
__device__ __forceinline__ int sk_bound(int p, int total, int P) { return (int)((long long)p * total / P); }

// Walks the segments (runs inside one tile) of a piece.
struct SkWalk {
  int pos, hi, num_k;
  int t, kb0, kb1; // current segment
  __device__ __forceinline__ void init(int lo, int hi_, int num_k_) {
    pos = lo;
    hi = hi_;
    num_k = num_k_;
  }
  __device__ __forceinline__ bool next() {
    if (pos >= hi) {
      return false;
    }
    t = pos / num_k;
    kb0 = pos - t * num_k;
    kb1 = min(num_k, kb0 + (hi - pos));
    pos += kb1 - kb0;
    return true;
  }
};

template <typename T>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(SK_THREADS, 1)
linear_sk_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 T *__restrict__ out, T const *__restrict__ bias, int M, int N, int K, int BN, int STAGES,
                 float4 *__restrict__ ws, int *__restrict__ cnt) {
  constexpr uint32_t A_BYTES = SK_BM * SK_BK * 2;
  uint32_t const B_BYTES = (uint32_t)(BN / 2) * SK_BK * 2; // a multiple of 1024 (BN % 16 == 0)
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[SK_MAX_STAGES], empty_bar[SK_MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ __align__(8) uint64_t part_bar[4][4]; // partial of epilogue warp q, 64-column chunk c has landed
  __shared__ uint32_t tmem_base_smem;
  // 128B-swizzled tiles need 1024-byte alignment; both CTAs compute the same offsets (the MMA addresses the peer's
  // shared memory with the leader's descriptors)
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sA = smem, *sB = smem + STAGES * A_BYTES;
  uint8_t *stage_out = sB + STAGES * B_BYTES; // 4 epilogue warps x 4 KB

  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  uint32_t const rank = sk_ctarank(); // 0 = leader
  int const num_k = K / SK_BK;
  int const m_tiles = (M + 255) / 256, n_tiles = (N + BN - 1) / BN, tiles = m_tiles * n_tiles;
  int const pair = blockIdx.x >> 1, P = gridDim.x >> 1;
  int const total = tiles * num_k;
  int const lo = sk_bound(pair, total, P), hi = sk_bound(pair + 1, total, P);

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
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer (both CTAs): this CTA's half of A and of W, straight through the segment boundaries =====
    int t = lo / num_k, kb = lo - t * num_k;
    int m_blk = t % m_tiles, n_blk = t / m_tiles;
    int s = 0;
    uint32_t ph = 1;
    // Launched as a programmatic dependent of the preceding kernel: the weights do not depend on it, so the W halves of
    // the first ring fill are requested at once; the activations wait for it (griddepcontrol.wait).
    int const pre = min(STAGES, hi - lo);
    {
      int pt = t, pkb = kb, pn = n_blk;
      for (int i = 0; i < pre; i++) {
        uint32_t const leader_full = sk_mapa(smem_u32(&full_bar[i]), 0);
        if (rank == 0) {
          mbar_expect_tx(&full_bar[i], 2 * (A_BYTES + B_BYTES));
        }
        sk_tma_load(smem_u32(sB + i * B_BYTES), &tmB, pkb * SK_BK, pn * BN + (int)rank * (BN / 2), leader_full);
        if (++pkb == num_k) {
          pkb = 0;
          pt++;
          pn = pt / m_tiles;
        }
      }
    }
    pdl_wait();
    pdl_trigger();
    for (int pos = lo; pos < hi; pos++) {
      bool const first_fill = pos - lo < pre; // (the slots of the first fill start empty; their W halves are on the way)
      if (!first_fill) {
        mbar_wait(&empty_bar[s], ph);
      }
      uint32_t const leader_full = sk_mapa(smem_u32(&full_bar[s]), 0);
      if (rank == 0 && !first_fill) {
        mbar_expect_tx(&full_bar[s], 2 * (A_BYTES + B_BYTES));
      }
      sk_tma_load(smem_u32(sA + s * A_BYTES), &tmA, kb * SK_BK, m_blk * 256 + (int)rank * SK_BM, leader_full);
      if (!first_fill) {
        sk_tma_load(smem_u32(sB + s * B_BYTES), &tmB, kb * SK_BK, n_blk * BN + (int)rank * (BN / 2), leader_full);
      }
      if (++s == STAGES) {
        s = 0;
        ph ^= 1u;
      }
      if (++kb == num_k) {
        kb = 0;
        t++;
        m_blk = t % m_tiles;
        n_blk = t / m_tiles;
      }
    }
  } else if (warp == 1 && lane == 0) {
    if (rank == 0) {
      // ===== MMA issuer (leader CTA only): one accumulator buffer per segment =====
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      uint32_t const idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);
      SkWalk w;
      w.init(lo, hi, num_k);
      int s = 0, i = 0;
      uint32_t ph = 0;
      while (w.next()) {
        int const b = i & 1;
        mbar_wait(&tmem_empty_bar[b], ((i >> 1) & 1) ^ 1); // both CTAs' epilogues have drained this buffer
        tc_fence_after();
        uint32_t const tmem_d = tmem_base + (uint32_t)b * 256u;
        for (int kb = w.kb0; kb < w.kb1; kb++) {
          mbar_wait(&full_bar[s], ph);
          tc_fence_after();
          uint64_t const da = umma_desc_sw128(smem_u32(sA + s * A_BYTES));
          uint64_t const db = umma_desc_sw128(smem_u32(sB + s * B_BYTES));
#pragma unroll
          for (int k = 0; k < SK_BK / 16; k++) {
            sk_mma(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, ((kb - w.kb0) | k) != 0);
          }
          sk_commit(&empty_bar[s]); // slot reusable in BOTH CTAs once these MMAs have read it
          if (++s == STAGES) {
            s = 0;
            ph ^= 1u;
          }
        }
        sk_commit(&tmem_full_bar[b]);
        i++;
      }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs): this CTA's 128 x BN half of each segment's accumulator =====
    int const q = warp & 3, wid = (int)rank * 4 + q; // this warp may touch TMEM lanes [32q, 32q+32)
    int const nchunk = (BN + 63) >> 6;
    uint8_t *const part = smem + (size_t)q * (BN >> 4) * 2048;
    uint8_t *const stg = stage_out + (warp - 4) * 4096;
    SkWalk w;
    w.init(lo, hi, num_k);
    int i = 0;
    uint32_t part_ph = 0;
    while (w.next()) {
      int const m_blk = w.t % m_tiles, n_blk = w.t / m_tiles;
      int const b = i & 1;
      mbar_wait(&tmem_full_bar[b], (i >> 1) & 1);
      tc_fence_after();
      int const row = m_blk * 256 + (int)rank * SK_BM + q * 32 + lane;
      uint32_t const tsrc = tmem_base + (uint32_t)b * 256u + ((uint32_t)(q * 32) << 16);
      uint32_t ra[32], rb[32];
      if (w.kb0 > 0) {
        // ----- not the head of its tile: park the fp32 partial in this pair's workspace slot (per 16 columns:
        // 4 x 32 lanes x float4, so that these stores and the owner's bulk copies are contiguous), raise the flag -----
        float4 *const my = ws + ((size_t)(pair * 8 + wid) * (BN >> 4)) * 128 + lane;
        tmem_ld32_nowait(tsrc, ra);
        tmem_ld32_nowait(tsrc + 32u, rb);
#pragma unroll 1
        for (int c = 0; c < nchunk; c++) {
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
          if (c + 1 < nchunk) {
            tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 64), ra);
            tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 96), rb);
          }
        }
        __threadfence();
        __syncwarp();
        if (lane == 0) {
          atomicExch(&cnt[pair * 8 + wid], 1);
        }
      } else {
        if (w.kb1 < num_k) {
          // ----- head of a tile that runs past this pair's piece (its LAST segment: the operand ring is idle, every
          // MMA has completed): add the partials of the following pairs, in pair order, into the accumulator -----
          int const tile_end = (w.t + 1) * num_k;
          for (int pq = pair + 1; pq < P && sk_bound(pq, total, P) < tile_end; pq++) {
            if (lane == 0) {
              int *const theirs = cnt + pq * 8 + wid;
              int seen;
              do {
                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(theirs) : "memory");
              } while (seen == 0);
              *theirs = 0; // ready for the next launch on this stream (the writer raises its flag once per launch)
              asm volatile("fence.proxy.async;" ::: "memory"); // the partials were written through the generic proxy
              float4 const *const src = ws + ((size_t)(pq * 8 + wid) * (BN >> 4)) * 128;
              for (int c = 0; c < nchunk; c++) {
                uint32_t const bytes = (uint32_t)(min(64, BN - c * 64) >> 4) * 2048u;
                mbar_expect_tx(&part_bar[q][c], bytes);
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                             ::"r"(smem_u32(part + (size_t)c * 8192)), "l"(src + (size_t)c * 512), "r"(bytes),
                               "r"(smem_u32(&part_bar[q][c])) : "memory");
              }
            }
            __syncwarp();
#pragma unroll 1
            for (int c = 0; c < nchunk; c++) {
              int const c0 = c * 64;
              tmem_ld32_nowait(tsrc + (uint32_t)c0, ra);
              tmem_ld32_nowait(tsrc + (uint32_t)(c0 + 32), rb);
              tmem_ld_wait32(ra);
              reg_tie32(rb);
              mbar_wait(&part_bar[q][c], part_ph);
              float4 const *p4 = reinterpret_cast<float4 const *>(part + (size_t)c * 8192) + lane;
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
              tmem_st32(tsrc + (uint32_t)c0, ra);
              tmem_st32(tsrc + (uint32_t)(c0 + 32), rb);
            }
            tmem_st_wait();
            __syncwarp(); // every lane has read this piece: the next one may overwrite it
            part_ph ^= 1u;
          }
        }
        // ----- finish: 64-column chunks TMEM -> registers (+ bias, convert) -> this warp's 32 x 128 B staging tile
        // (16-byte units XOR-swizzled by row) -> global with 8 lanes per row (every warp store covers 4 full lines) -----
        int const col0 = n_blk * BN, row0 = row - lane;
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
      }
      // hand the buffer back to the leader's MMA issuer
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        uint32_t const remote = sk_mapa(smem_u32(&tmem_empty_bar[b]), 0);
        asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(remote) : "memory");
      }
      i++;
    }
  }
  tc_fence_before();
  // neither CTA frees tensor memory / exits while the peer may still use its smem or TMEM; execution ordering only
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.aligned;" ::: "memory");
  if (warp == 2) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

template <typename T>
int launch_sk(void const *in, void *out, void const *w, void const *bias, int dtype, int M, int K, int N, int BN, int P,
              void *ws, int *cnt, cudaStream_t s) {
  size_t const stage = (size_t)(SK_BM + BN / 2) * SK_BK * 2;
  int const stages = std::min<int>(SK_MAX_STAGES, (int)(((size_t)208 << 10) / stage));
  CUtensorMap tmA, tmB;
  int rc = make_tensor_map_2d(&tmA, dtype, in, M, K, SK_BM);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(&tmB, dtype, w, N, K, BN / 2);
  if (rc != FFB_OK) {
    return rc;
  }
  FFB_SET_MAX_SMEM_ONCE(linear_sk_kernel<T>, ((size_t)225 << 10) + 1024);
  size_t const smem = (size_t)stages * stage + 16384 + 1024; // ring + epilogue staging + alignment slack
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(2 * P);
  cfg.blockDim = dim3(SK_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = gemm_pdl_enabled() ? 1 : 0;
  FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, linear_sk_kernel<T>, tmA, tmB, (T *)out, (T const *)bias, M, N, K, BN, stages,
                                  (float4 *)ws, cnt));
  return FFB_OK;
}

} // namespace

// Tile width: N cut into ceil(N / 256) equal tiles, rounded up to the MMA's multiple of 16.
static int sk_tile_width(int N) {
  int const nt = (N + 255) / 256;
  return std::min(256, ((N + nt - 1) / nt + 15) / 16 * 16);
}

// Used when the plain persistent schedule would leave a partial last wave and every pair still gets a few k-blocks;
// the merge needs the split-K workspace of the handle (one slot and 8 flags per pair).  FFB_SK=0 disables, =1 forces.
int linear_sk_supported(int dtype, int M, int K, int N, void const *ws, int const *cnt) {
  char const *e = ffb::env("FFB_SK");
  if (e && e[0] == '0') {
    return 0;
  }
  if (!(dtype == FFB_DT_HALF || dtype == FFB_DT_BF16) || M <= 128 || K % SK_BK != 0 || N % 16 != 0 || !ws || !cnt) {
    return 0;
  }
  int const pairs = kNumSMs / 2, num_k = K / SK_BK, bn = sk_tile_width(N);
  long const tiles = (long)((M + 255) / 256) * ((N + bn - 1) / bn);
  if (tiles * num_k < (long)pairs * 8 || tiles * num_k > (1L << 30)) {
    return 0;
  }
  if (e && e[0] == '1') {
    return 1;
  }
  // Measured (profiles/README.md, M = 256 ... 1024 at the 7B shapes): the cut pays when the plain schedule's last wave
  // is less than half full (M = 512, N = 12288: 96 tiles on 74 pairs, 49.3 -> 46.3 us); a fuller last wave, or fewer
  // tiles than pairs (every tile cut into several pieces, each merge a trip through L2), is faster uncut.
  long const rem = tiles % pairs;
  return tiles > pairs && rem != 0 && 2 * rem < pairs;
}

int launch_linear_sk(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                     void *ws, int *cnt, cudaStream_t s) {
  int const bn = sk_tile_width(N);
  int P = kNumSMs / 2;
  if (char const *e = ffb::env("FFB_SK_PAIRS")) {
    int const v = atoi(e);
    if (v >= 1 && v <= kNumSMs / 2) {
      P = v;
    }
  }
  if (dtype == FFB_DT_HALF) {
    return launch_sk<__half>(in, out, w, bias, dtype, M, K, N, bn, P, ws, cnt, s);
  }
  return launch_sk<__nv_bfloat16>(in, out, w, bias, dtype, M, K, N, bn, P, ws, cnt, s);
}

} // namespace ffb
