// linear_gemv.cu — persistent weight-streaming projection for decode batches (M <= 16 tokens, DT weights).
// Replaces the cublasGemmEx call of Kernels::Linear::forward_kernel (src/ops/kernels/linear_kernels.cu:507-528)
// for the case where the GEMM is a pure stream over the weights (7B decode: 134 MB of qkv+o weights per layer
// against 2 GFLOP).  One CTA per SM stays resident for the whole launch:
//   * 16 consumer warps each own a fixed K/16 slice of the reduction dimension and keep the matching
//     activation fragments ([16 tokens] x [K/16]) in REGISTERS for the whole kernel (loaded once through the
//     ring with ldmatrix), so shared memory holds nothing but the weight stream;
//   * a producer thread streams weight tiles (8 output features x K, one cp.async.bulk per row) through a
//     ~200 KiB mbarrier ring; the stream runs on across tile boundaries, so the memory system never drains
//     until the CTA's last tile;
//   * per tile each warp issues K/256 ldmatrix.x4 + K/16/16 mma.sync m16n8k16 (tokens = M operand, fp32
//     accumulate) and the 16 partial sums are merged in warp order (deterministic).
// 8-feature tiles keep the static round-robin schedule within one tile (64 KiB at K = 4096) of perfect balance.
//
// FUSED o_proj + AllReduce (FUSED = true; replaces Linear -> AllReduce, src/ops/fused.cu:448-495 +
// src/parallel_ops/kernels/allreduce_kernels.cu:64-73, for decode batches): instead of storing its partial sums, a
// CTA PUSHES them over NVLink into slot [my rank] of every rank's receive area as 8-byte {fp32 value, epoch} words
// (one store each: the epoch travelling WITH the value is the arrival flag, so there is no fence, no separate flag
// and only one NVLink crossing on the critical path), then polls its own slots until every peer's words carry
// this epoch and sums them in rank order (deterministic, fp32) into the final output.  CTA i of every rank owns the
// same output tiles, so the exchange needs no grid-wide barrier and overlaps with the weight stream of the slower
// CTAs.  Receive areas are double buffered by epoch parity: a rank can only have pushed epoch e+1 after it finished
// reading epoch e (stream order), so by the time this rank has read everybody's e+1 it may overwrite the e slots.
#include <cstdlib>

#include "peer.h"
#include "tc_common.cuh"

namespace ffb {
namespace {

constexpr int GV_CONSUMERS = 16, GV_THREADS = (GV_CONSUMERS + 1) * 32;
constexpr int GV_BN = 8;
constexpr int GV_RING_BYTES = 200 * 1024;

__device__ __forceinline__ void gv_ldsm4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
template <typename T>
__device__ __forceinline__ void gv_mma(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void gv_mma<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void gv_mma<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// KSTEPS = K / 16 / 16: mma k-steps per warp per tile (2, 4, 8 or 16 for K = 512, 1024, 2048, 4096)
template <typename T, int KSTEPS, bool FUSED>
__global__ void __launch_bounds__(GV_THREADS, 1)
linear_gemv_kernel(T const *__restrict__ A, T const *__restrict__ W, T const *__restrict__ bias, T *__restrict__ out,
                   int M, int N, const __grid_constant__ PeerTable pt) {
  constexpr int K = KSTEPS * 16 * GV_CONSUMERS;
  constexpr int WROW = K * 2 + 16; // padded row (bytes): 8 rows hit 8 distinct bank groups in ldmatrix
  constexpr int STAGE_BYTES = GV_BN * WROW;
  constexpr int STAGES = GV_RING_BYTES / STAGE_BYTES; // 3 at K = 4096
  static_assert(STAGES >= 3, "ring too small");
  extern __shared__ __align__(16) uint8_t gv_smem[];
  uint8_t *ring = gv_smem;
  float *red = reinterpret_cast<float *>(ring + STAGES * STAGE_BYTES); // [2][16 warps][16][8]
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES];

  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const epoch = FUSED ? *reinterpret_cast<int const volatile *>(pt.epoch) + 1 : 0; // this collective's number
  int const ntiles = N / GV_BN;
  int const my_tiles = (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

  if (tid == 0) {
    for (int s = 0; s < STAGES; s++) {
      tc::mbar_init(&full_bar[s], 1);
      tc::mbar_init(&empty_bar[s], GV_CONSUMERS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  // The activations travel through ring stages 0 and 1 first (rows 0-7 and 8-15); rows past M are zero.
  for (int i = tid; i < (16 - M) * (WROW / 16); i += GV_THREADS) {
    reinterpret_cast<uint4 *>(ring + (size_t)M * WROW)[i] = make_uint4(0, 0, 0, 0);
  }
  __syncthreads();

  if (warp == GV_CONSUMERS) {
    // ===== producer =====
    if (lane == 0) {
      // it = 0, 1: activation halves; it >= 2: weight tiles.  Stage 2 (first weight tile) is requested
      // before the activations so the HBM stream starts immediately.
      int const m0 = M < 8 ? M : 8, m1 = M - m0;
      if (my_tiles > 0) {
        int const n0 = (int)blockIdx.x * GV_BN;
        tc::mbar_expect_tx(&full_bar[2], GV_BN * K * 2);
#pragma unroll
        for (int r = 0; r < GV_BN; r++) {
          tc::bulk_load_1d(ring + 2 * STAGE_BYTES + r * WROW, W + (size_t)(n0 + r) * K, K * 2, &full_bar[2]);
        }
      }
      tc::mbar_expect_tx(&full_bar[0], (uint32_t)(m0 * K * 2));
      for (int r = 0; r < m0; r++) {
        tc::bulk_load_1d(ring + (size_t)r * WROW, A + (size_t)r * K, K * 2, &full_bar[0]);
      }
      tc::mbar_expect_tx(&full_bar[1], (uint32_t)(m1 * K * 2));
      for (int r = 0; r < m1; r++) {
        tc::bulk_load_1d(ring + (size_t)(8 + r) * WROW, A + (size_t)(8 + r) * K, K * 2, &full_bar[1]);
      }
      for (int t = 1; t < my_tiles; t++) {
        int const it = t + 2, s = it % STAGES;
        int const n0 = ((int)blockIdx.x + t * (int)gridDim.x) * GV_BN;
        tc::mbar_wait(&empty_bar[s], ((it / STAGES) & 1) ^ 1);
        tc::mbar_expect_tx(&full_bar[s], GV_BN * K * 2);
#pragma unroll
        for (int r = 0; r < GV_BN; r++) {
          tc::bulk_load_1d(ring + s * STAGE_BYTES + r * WROW, W + (size_t)(n0 + r) * K, K * 2, &full_bar[s]);
        }
      }
    }
  } else {
    // ===== consumers: warp w owns k in [w*K/16, (w+1)*K/16) =====
    uint32_t const ring_addr = tc::smem_u32(ring);
    int const g = lane >> 2, tq = lane & 3;
    int const kw = warp * (KSTEPS * 16);
    uint32_t a[KSTEPS][4];
    {
      tc::mbar_wait(&full_bar[0], 0);
      tc::mbar_wait(&full_bar[1], 0);
      int const a_row = (lane & 7) + ((lane >> 3) & 1) * 8, a_kc = (lane >> 4) * 8;
      uint32_t const abase = ring_addr + a_row * WROW + (kw + a_kc) * 2;
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ks++) {
        gv_ldsm4(abase + ks * 32, a[ks][0], a[ks][1], a[ks][2], a[ks][3]);
      }
      __syncwarp();
      if (lane == 0) {
        tc::mbar_arrive(&empty_bar[0]);
        tc::mbar_arrive(&empty_bar[1]);
      }
    }
    int const b_row = lane & 7, b_kc = (lane >> 3) * 8;
    for (int t = 0; t < my_tiles; t++) {
      int const it = t + 2, s = it % STAGES;
      int const n0 = ((int)blockIdx.x + t * (int)gridDim.x) * GV_BN;
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      tc::mbar_wait(&full_bar[s], (it / STAGES) & 1);
      uint32_t const wbase = ring_addr + s * STAGE_BYTES + b_row * WROW + (kw + b_kc) * 2;
      uint32_t b[KSTEPS / 2][4];
#pragma unroll
      for (int j = 0; j < KSTEPS / 2; j++) { // 32 k per ldmatrix.x4: (b0,b1) of k-step 2j and of 2j+1
        gv_ldsm4(wbase + j * 64, b[j][0], b[j][1], b[j][2], b[j][3]);
      }
      __syncwarp();
      if (lane == 0) {
        tc::mbar_arrive(&empty_bar[s]); // the tile now lives in registers
      }
#pragma unroll
      for (int ks = 0; ks < KSTEPS; ks++) {
        gv_mma<T>(acc, a[ks][0], a[ks][1], a[ks][2], a[ks][3], b[ks >> 1][(ks & 1) * 2], b[ks >> 1][(ks & 1) * 2 + 1]);
      }
      // merge the 16 k-slices in warp order
      float *slot = red + (size_t)(t & 1) * (GV_CONSUMERS * 128);
      slot[warp * 128 + g * 8 + 2 * tq] = acc[0];
      slot[warp * 128 + g * 8 + 2 * tq + 1] = acc[1];
      slot[warp * 128 + (g + 8) * 8 + 2 * tq] = acc[2];
      slot[warp * 128 + (g + 8) * 8 + 2 * tq + 1] = acc[3];
      asm volatile("bar.sync 1, %0;" ::"n"(GV_CONSUMERS * 32) : "memory");
      if (tid < 128) {
        int const r = tid >> 3, c = tid & 7;
        float v = 0.f;
#pragma unroll
        for (int w = 0; w < GV_CONSUMERS; w++) {
          v += slot[w * 128 + tid];
        }
        if (FUSED) {
          if (r < M) { // {value, epoch} into this rank's slot of every rank's receive area (8 features = 64 B runs)
            size_t const off = ((size_t)((epoch & 1) * pt.nranks + pt.rank) * kPeerMaxTokens + r) * pt.max_n + n0 + c;
            for (int p = 0; p < pt.nranks; p++) {
              // one 64-bit scalar store: single-copy atomic, so the reader never sees a value without its epoch
              unsigned long long const word = ((unsigned long long)(uint32_t)epoch << 32) | __float_as_uint(v);
              asm volatile("st.relaxed.sys.global.b64 [%0], %1;" ::"l"(pt.recv[p] + 2 * off), "l"(word) : "memory");
            }
          }
        } else if (r < M) {
          if (bias) {
            v += DT<T>::to_f(bias[n0 + c]);
          }
          out[(size_t)r * N + n0 + c] = DT<T>::from_f(v);
        }
      }
    }
    if (FUSED) {
      // ---- reduce in rank order; a word has arrived when its epoch field matches ----
      float const *mine = pt.recv[pt.rank] + (size_t)2 * (epoch & 1) * pt.nranks * kPeerMaxTokens * pt.max_n;
      for (int i = tid; i < my_tiles * 128; i += GV_CONSUMERS * 32) {
        int const t = i >> 7, r = (i >> 3) & 15, c = i & 7;
        int const n0 = ((int)blockIdx.x + t * (int)gridDim.x) * GV_BN;
        if (r < M) {
          // the reference adds the bias on EVERY shard before the sum (linear_kernels.cu:531, then allreduce): nranks x bias
          float v = bias ? (float)pt.nranks * DT<T>::to_f(bias[n0 + c]) : 0.f;
          for (int p = 0; p < pt.nranks; p++) {
            float const *src = mine + 2 * (((size_t)p * kPeerMaxTokens + r) * pt.max_n + n0 + c);
            unsigned long long word;
            do {
              asm volatile("ld.relaxed.sys.global.b64 %0, [%1];" : "=l"(word) : "l"(src) : "memory");
            } while ((int)(word >> 32) != epoch);
            v += __uint_as_float((uint32_t)word);
          }
          out[(size_t)r * N + n0 + c] = DT<T>::from_f(v);
        }
      }
      // the last CTA to finish closes the epoch (every CTA read it at kernel start)
      asm volatile("bar.sync 1, %0;" ::"n"(GV_CONSUMERS * 32) : "memory");
      if (tid == 0) {
        int const prev = atomicAdd(pt.done, 1);
        if (prev == (int)gridDim.x - 1) {
          *pt.done = 0;
          __threadfence();
          *reinterpret_cast<int volatile *>(pt.epoch) = epoch;
        }
      }
    }
  }
}

} // namespace

int linear_gemv_supported(int quant, int M, int K, int N) {
  if (ffb::env("FFB_DISABLE_GEMV")) {
    return 0;
  }
  return quant == FFB_DT_NONE && M >= 1 && M <= 16 && (K == 512 || K == 1024 || K == 2048 || K == 4096) && N % GV_BN == 0;
}

template <typename T, int KSTEPS, bool FUSED>
static int launch_gemv_t(void const *in, void *out, void const *w, void const *bias, int M, int N, cudaStream_t s,
                         PeerTable const &pt) {
  constexpr int K = KSTEPS * 16 * GV_CONSUMERS;
  constexpr int STAGE_BYTES = GV_BN * (K * 2 + 16);
  constexpr size_t smem = (size_t)(GV_RING_BYTES / STAGE_BYTES) * STAGE_BYTES + 2 * GV_CONSUMERS * 128 * 4;
  FFB_SET_MAX_SMEM_ONCE((linear_gemv_kernel<T, KSTEPS, FUSED>), smem);
  int const ntiles = N / GV_BN;
  int const grid = ntiles < kNumSMs ? ntiles : kNumSMs;
  linear_gemv_kernel<T, KSTEPS, FUSED><<<grid, GV_THREADS, smem, s>>>((T const *)in, (T const *)w, (T const *)bias, (T *)out, M, N, pt);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

template <bool FUSED>
static int launch_gemv_dispatch(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                                cudaStream_t s, PeerTable const &pt) {
  int const ks = K / (16 * GV_CONSUMERS);
  if (ks == 2) {
    return dtype == FFB_DT_HALF ? launch_gemv_t<__half, 2, FUSED>(in, out, w, bias, M, N, s, pt)
                                : launch_gemv_t<__nv_bfloat16, 2, FUSED>(in, out, w, bias, M, N, s, pt);
  }
  if (dtype == FFB_DT_HALF) {
    return ks == 4 ? launch_gemv_t<__half, 4, FUSED>(in, out, w, bias, M, N, s, pt)
                   : (ks == 8 ? launch_gemv_t<__half, 8, FUSED>(in, out, w, bias, M, N, s, pt)
                              : launch_gemv_t<__half, 16, FUSED>(in, out, w, bias, M, N, s, pt));
  }
  return ks == 4 ? launch_gemv_t<__nv_bfloat16, 4, FUSED>(in, out, w, bias, M, N, s, pt)
                 : (ks == 8 ? launch_gemv_t<__nv_bfloat16, 8, FUSED>(in, out, w, bias, M, N, s, pt)
                            : launch_gemv_t<__nv_bfloat16, 16, FUSED>(in, out, w, bias, M, N, s, pt));
}

int launch_linear_gemv(int dtype, void const *in, void *out, void const *w, void const *bias, int M, int K, int N,
                       cudaStream_t s) {
  PeerTable none;
  return launch_gemv_dispatch<false>(dtype, in, out, w, bias, M, K, N, s, none);
}

// o_proj + AllReduce in one launch.  Every rank of the communicator must call it with the same M, K, N.
int linear_gemv_allreduce_supported(PeerTable const *pt, int quant, int M, int K, int N) {
  return pt && pt->nranks >= 2 && M <= kPeerMaxTokens && N <= pt->max_n && (N / GV_BN < kNumSMs ? N / GV_BN : kNumSMs) <= kPeerMaxCtas &&
         linear_gemv_supported(quant, M, K, N);
}

int launch_linear_gemv_allreduce(PeerTable const *pt, int dtype, void const *in, void *out, void const *w, void const *bias,
                                 int M, int K, int N, cudaStream_t s) {
  return launch_gemv_dispatch<true>(dtype, in, out, w, bias, M, K, N, s, *pt);
}

} // namespace ffb
