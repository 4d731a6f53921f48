// linear_qtc.cu — projection over QUANTISED weights for decode-sized batches (<= 64 tokens, half activations): the
// int8 / int4 blob ([values | offsets | scales], src/ops/kernels/decompress_kernels.cu:27-70) is streamed once by TMA,
// dequantised in shared memory with the bit-exact rule of dequant.cuh, and multiplied on the 5th-generation tensor
// cores with the WEIGHTS as the 128-row operand: D[128 features][tokens] += Wq[128][k] x X[tokens][k]^T.
// Replaces, for quantised layers, the H2D copy + decompress pass + cublasGemmEx of Kernels::Linear::forward_kernel
// (src/ops/kernels/linear_kernels.cu:464-528).
//
// Why a third decode kernel: linear_stream / linear_qgemv feed mma.sync from registers.  ncu on linear_qgemv (7B
// qkv_proj, int8, 16 tokens, 24 us for 57 MB): 6.1 instructions per weight element of which 3.5 are the dequantisation
// itself, 128 registers per thread (activation fragments of the whole row + accumulators) and so 4 warps per
// scheduler, IPC 1.8.  With tcgen05 the accumulator lives in tensor memory and the activations in shared memory; the
// converter warps hold nothing but the elements they convert (16 warps, ~48 registers), the MMA costs one thread.
//
// Work split: the flat space of (128-feature tile, 128-wide k-step) pairs, tile-major, is cut into one contiguous range
// per CTA (one CTA per SM): tail of a tile, whole tiles, head of a tile.  A cut tile is finished by the CTA that holds
// its head - the last segment of that CTA - from the fp32 partials (tokens x 128 floats) that the following CTAs parked
// for their first segments, added in CTA order (deterministic).  Per k-step:
//   warp 16 (one lane)  TMA: 128 x 128 codes (uint8 tensor map, no swizzle), offsets / scales of the tile's four
//                       quantisation groups, the tokens' k-step of activations (two 128B-swizzled boxes)
//   warps 0-15          converters: thread (row set, 8 consecutive k) dequantises 8 k x 4 rows of ONE group with the
//                       reciprocals of its own 8 columns (no table, so no CTA-wide barrier: the warps run ahead of
//                       each other), and writes the fp16 result as the 128B-swizzled K-major A operand
//   warp 17 (one lane)  8 x tcgen05.mma 128 x Npad x 16; tcgen05.commit releases the operand buffer and the stage
//   warps 0-3           at the end of a segment: tensor memory -> registers -> out / partial
#include <cuda.h>

#include <algorithm>
#include <cstdio>
#include <cstring>

#include "dequant.cuh"
#include "tc_common.cuh"

namespace ffb {
using namespace tc;
// linear_tc.cu
int make_tensor_map_u8(CUtensorMap *tm, void const *ptr, long rows, long row_bytes, int box_rows, int box_bytes);
int make_tensor_map_nd(CUtensorMap *tm, int elem, bool is_half, int rank, void const *ptr, unsigned long long const *dims,
                       unsigned long long const *strides, unsigned const *box, bool swizzle128);

namespace {

constexpr int QT_TM = 128;              // features per tile (four quantisation groups)
constexpr int QT_KS = 128;              // k per step (two 64-wide swizzle atoms)
constexpr int QT_CONV_WARPS = 16, QT_CONV_THREADS = QT_CONV_WARPS * 32;
constexpr int QT_PRODUCER_WARP = 16, QT_MMA_WARP = 17, QT_THREADS = 18 * 32;
constexpr int QT_MAX_STAGES = 8;
#ifndef QT_NB
#define QT_NB 4
#endif
constexpr int QT_NBUF = QT_NB; // converted operand buffers (a power of two): the converters may run NBUF - 1 k-steps ahead of the MMA
constexpr uint32_t QT_ABUF_BYTES = QT_TM * QT_KS * 2; // one converted operand buffer: two [128][64] fp16 atoms
constexpr uint32_t QT_META_BYTES = 2 * 4 * QT_KS * 2; // offsets then scales of 4 groups x 128 k

#if defined(FFB_QT_TRACE_BUILD) && defined(QT_TRACE_ROLES)
// nvcc -DFFB_QT_TRACE_BUILD -DQT_TRACE_ROLES: CTA `QT_TRACE_CTA` stamps %globaltimer (event id, time) pairs; the host prints them every 8th launch
__device__ long long g_qt_cta[160][3];
__device__ long long g_qt_trace[4][64][2];
__device__ int g_qt_n[4];
#define QT_TR(role, ev)                                                              \
  do {                                                                               \
    if (blockIdx.x == 70 && g_qt_n[role] < 64) {                                     \
      long long t_;                                                                  \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                         \
      g_qt_trace[role][g_qt_n[role]][0] = ev;                                        \
      g_qt_trace[role][g_qt_n[role]][1] = t_;                                        \
      g_qt_n[role]++;                                                                \
    }                                                                                \
  } while (0)
#else
#ifdef FFB_QT_TRACE_BUILD
__device__ long long g_qt_cta[160][3];
__device__ long long g_qt_trace[4][64][2];
__device__ int g_qt_n[4];
#endif
#define QT_TR(role, ev) ((void)0)
#endif

// segments (runs inside one tile) of a CTA's range of the flat (tile, k-step) space
struct QtWalk {
  int pos, hi, spt;
  int t, ks0, ks1;
  __device__ __forceinline__ void init(int lo, int hi_, int spt_) {
    pos = lo;
    hi = hi_;
    spt = spt_;
  }
  __device__ __forceinline__ bool next() {
    if (pos >= hi) {
      return false;
    }
    t = pos / spt;
    ks0 = pos - t * spt;
    ks1 = min(spt, ks0 + (hi - pos));
    pos += ks1 - ks0;
    return true;
  }
};
__device__ __forceinline__ int qt_bound(int c, int total, int G) { return (int)((long long)c * total / G); }

// element-wise rule for a k-step with a degenerate scale (0, inf, NaN, negative): out of line, off the fast path
// codes: first code byte of this k-step in row 0 (row pitch `pitch`); offs / scales: this thread's 8 columns
template <int QB>
__device__ __noinline__ void qt_slow_convert(uint8_t const *codes, int pitch, __half const *offs, __half const *scales,
                                             uint8_t *abuf, int g, int j, int c) {
  for (int i = 0; i < 4; i++) {
    int const r = 32 * g + j + 8 * i;
    uint8_t const *src = codes + r * pitch + (QB == 8 ? 8 * c : 4 * c);
    __half h[8];
    for (int e = 0; e < 8; e++) {
      int q;
      if (QB == 8) {
        q = src[e];
      } else {
        int const byte = src[e >> 1];
        q = (e & 1) ? (byte & 0xF) : (byte >> 4);
      }
      h[e] = dq_elem(q, dq_coef(scales[e], offs[e]));
    }
    uint4 v;
    v.x = (uint32_t)__half_as_ushort(h[0]) | ((uint32_t)__half_as_ushort(h[1]) << 16);
    v.y = (uint32_t)__half_as_ushort(h[2]) | ((uint32_t)__half_as_ushort(h[3]) << 16);
    v.z = (uint32_t)__half_as_ushort(h[4]) | ((uint32_t)__half_as_ushort(h[5]) << 16);
    v.w = (uint32_t)__half_as_ushort(h[6]) | ((uint32_t)__half_as_ushort(h[7]) << 16);
    *reinterpret_cast<uint4 *>(abuf + (c >> 3) * (QT_TM * 128) + r * 128 + (((c & 7) ^ (r & 7)) << 4)) = v;
  }
}

// SUBS = 128-wide k-steps per ring stage: a stage's codes are ONE box of 128 rows x SUBS * 128 k, so that every row
// visit reads 256 contiguous bytes (a 128-byte visit per row and k-step is one DRAM activate per 128 bytes: ~2 TB/s)
template <int QB, int SUBS>
__global__ void __launch_bounds__(QT_THREADS, 1)
linear_qtc_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmX,
                  const __grid_constant__ CUtensorMap tmM, uint8_t const *__restrict__ blob, __half const *__restrict__ bias, __half *__restrict__ out,
                  float *__restrict__ ws, int *__restrict__ cnt, int M, int N, int K, int NPAD, int STAGES) {
  constexpr int ROWB = QT_KS * QB / 8;               // code bytes of one feature row per k-step
  constexpr int PITCH = SUBS * ROWB;                 // ... per stage
  constexpr uint32_t CODE_BYTES = QT_TM * PITCH;
  constexpr uint32_t META_BYTES = SUBS * QT_META_BYTES; // offsets [4][SUBS * 128], then scales
  uint32_t const act_bytes = (uint32_t)NPAD * 256u * SUBS; // 2 SUBS [NPAD][64] fp16 atoms
  uint32_t const stage_bytes = act_bytes + CODE_BYTES + META_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[QT_MAX_STAGES], empty_bar[QT_MAX_STAGES];
  __shared__ __align__(8) uint64_t conv_bar[QT_NBUF], afree_bar[QT_NBUF], tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_smem;
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *ring = smem;
  uint8_t *abuf = ring + (size_t)STAGES * stage_bytes;                       // NBUF x 32 KB, 1024-aligned

  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef FFB_QT_TRACE_BUILD
  if (tid == 0) {
    long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    g_qt_cta[blockIdx.x][0] = t_;
  }
#endif
  int const spt = K / (QT_KS * SUBS), tiles = N / QT_TM, total = tiles * spt; // work unit: one stage
  int const cta = blockIdx.x, G = gridDim.x;
  int const lo = qt_bound(cta, total, G), hi = qt_bound(cta + 1, total, G);
  size_t const nvalues = (size_t)N * K;
  __half const *offsets = reinterpret_cast<__half const *>(blob + (QB == 4 ? nvalues / 2 : nvalues));
  __half const *scales = offsets + nvalues / 32;

  if (warp == QT_PRODUCER_WARP && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmW) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmX) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&tmM) : "memory");
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], QT_CONV_WARPS + 1); // every converter warp + the MMA's commit
    }
    for (int b = 0; b < QT_NBUF; b++) {
      mbar_init(&conv_bar[b], QT_CONV_WARPS);
      mbar_init(&afree_bar[b], 1);
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == QT_MMA_WARP) { // two accumulators of up to 64 fp32 columns
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(128u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;

  if (warp == QT_PRODUCER_WARP) {
    if (lane == 0) {
      // ===== TMA producer =====
      int t = lo / spt, ks = lo - t * spt;
      int s = 0;
      uint32_t ph = 1;
      QT_TR(0, 0);
      for (int pos = lo; pos < hi; pos++) {
        mbar_wait_sleep(&empty_bar[s], ph);
        QT_TR(0, 100 + pos - lo);
        uint8_t *stage = ring + (size_t)s * stage_bytes;
        mbar_expect_tx(&full_bar[s], stage_bytes);
        int const k0 = ks * QT_KS * SUBS, n0 = t * QT_TM;
        // four copies per stage (a single thread issues them: every copy instruction costs ~0.1 us of its time):
        // codes, offsets and scales of the tile's four groups (rows g and N / 32 + g of the [2 N / 32][K] meta matrix),
        // and the tokens' activations as 2 SUBS swizzle atoms in one 3-D box
        tma_load_2d(stage + act_bytes, &tmW, k0 * QB / 8, n0, &full_bar[s]);
        uint8_t *meta = stage + act_bytes + CODE_BYTES;
        tma_load_2d(meta, &tmM, k0, n0 / 32, &full_bar[s]);
        tma_load_2d(meta + META_BYTES / 2, &tmM, k0, N / 32 + n0 / 32, &full_bar[s]);
        tma_load_3d(stage, &tmX, 0, 0, k0 / 64, &full_bar[s]);
        if (++s == STAGES) {
          s = 0;
          ph ^= 1u;
        }
        if (++ks == spt) {
          ks = 0;
          t++;
        }
      }
    }
  } else if (warp == QT_MMA_WARP) {
    if (lane == 0) {
      // ===== MMA issuer: D[128 features][NPAD tokens] (+)= A (converted weights) x B (activations) =====
      uint32_t const idesc = (1u << 4) | ((uint32_t)(NPAD >> 3) << 17) | ((uint32_t)(QT_TM >> 4) << 24); // F16 x F16 -> F32
      QtWalk w;
      w.init(lo, hi, spt);
      int i = 0, seg = 0, s = 0;
      uint32_t ph = 0;
      while (w.next()) {
        int const ab = seg & 1;
        mbar_wait_sleep(&tmem_empty_bar[ab], ((seg >> 1) & 1) ^ 1);
        tc_fence_after();
        uint32_t const tmem_d = tmem_base + (uint32_t)ab * 64u;
        for (int ks = w.ks0; ks < w.ks1; ks++) {
          mbar_wait_sleep(&full_bar[s], ph);                 // activations of this stage (TMA)
          uint32_t const x_base = smem_u32(ring + (size_t)s * stage_bytes);
#pragma unroll 1
          for (int ss = 0; ss < SUBS; ss++, i++) {
            int const b = i % QT_NBUF;
            mbar_wait_sleep(&conv_bar[b], (i / QT_NBUF) & 1); // converted weights of this k-step
            QT_TR(1, 100 + i);
            tc_fence_after();
            uint32_t const a_base = smem_u32(abuf + (size_t)b * QT_ABUF_BYTES);
#pragma unroll
            for (int a = 0; a < 2; a++) {
              uint64_t const da = umma_desc_sw128(a_base + a * (QT_TM * 128));
              uint64_t const db = umma_desc_sw128(x_base + (2 * ss + a) * (NPAD * 128));
#pragma unroll
              for (int k = 0; k < 4; k++) {
#ifdef QT_EXP_ONE_MMA // experiment build: one MMA per k-step instead of eight
                if (a + k > 0) {
                  continue;
                }
#endif
                tc_mma_f16(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc,
                           (ks > w.ks0 || ss > 0 || a > 0 || k > 0) ? 1u : 0u);
              }
            }
            tc_commit(&afree_bar[b]);
          }
          tc_commit(&empty_bar[s]);
          if (++s == STAGES) {
            s = 0;
            ph ^= 1u;
          }
        }
        tc_commit(&tmem_full_bar[ab]);
        seg++;
      }
    }
  } else {
    // ===== converters (and, warps 0-3, the epilogue of every segment) =====
    int const c = tid & 15, rs = tid >> 4, g = rs >> 3, j = rs & 7; // 8 k (chunk c) x rows 32 g + j + 8 i of group g
    uint32_t const code_off = (uint32_t)((32 * g + j) * PITCH + (QB == 8 ? 8 * c : 4 * c));
    uint32_t const dst_off = (uint32_t)((c >> 3) * (QT_TM * 128) + (32 * g + j) * 128 + (((c & 7) ^ j) << 4));
    uint32_t const meta_off = (uint32_t)((g * SUBS * QT_KS + 8 * c) * 2);
    QtWalk w;
    w.init(lo, hi, spt);
    int i = 0, seg = 0, s = 0;
    uint32_t ph = 0;
    while (w.next()) {
      for (int ks = w.ks0; ks < w.ks1; ks++) {
        if (tid == 0) QT_TR(2, 100 + i);
        mbar_wait_sleep(&full_bar[s], ph);
        if (tid == 0) QT_TR(2, 200 + i);
        uint8_t const *stage = ring + (size_t)s * stage_bytes;
#pragma unroll 1
        for (int ss = 0; ss < SUBS; ss++, i++) {
        int const b = i % QT_NBUF;
        uint8_t const *codes = stage + act_bytes + ss * ROWB, *meta = stage + act_bytes + CODE_BYTES + ss * (QT_KS * 2);
        // this thread's 8 columns of group g: reciprocals computed here (8 MUFU per 32 elements) - a shared table would
        // need a CTA-wide barrier per k-step, and the warps would then convert in lock-step (measured: 1.4 us per step)
        uint4 const sc8 = *reinterpret_cast<uint4 const *>(meta + META_BYTES / 2 + meta_off);
        uint4 const of = *reinterpret_cast<uint4 const *>(meta + meta_off);
        float r[8];
        {
          float2 const a = __half22float2(*reinterpret_cast<__half2 const *>(&sc8.x));
          float2 const b2 = __half22float2(*reinterpret_cast<__half2 const *>(&sc8.y));
          float2 const c2 = __half22float2(*reinterpret_cast<__half2 const *>(&sc8.z));
          float2 const d2 = __half22float2(*reinterpret_cast<__half2 const *>(&sc8.w));
          r[0] = rcp_approx_ftz(a.x); r[1] = rcp_approx_ftz(a.y); r[2] = rcp_approx_ftz(b2.x); r[3] = rcp_approx_ftz(b2.y);
          r[4] = rcp_approx_ftz(c2.x); r[5] = rcp_approx_ftz(c2.y); r[6] = rcp_approx_ftz(d2.x); r[7] = rcp_approx_ftz(d2.y);
        }
        bool bad = false;
#pragma unroll
        for (int e = 0; e < 8; e++) {
          bad = bad || !dq_rcp_ok(r[e]);
        }
        if (tid == 0) QT_TR(2, 500 + i + (bad ? 1000 : 0) + (__float_as_int(r[0] + r[7]) == 12345 ? 1 : 0));
        mbar_wait_sleep(&afree_bar[b], ((i / QT_NBUF) & 1) ^ 1); // the MMAs that read this operand buffer NBUF steps ago are done
        if (tid == 0) QT_TR(2, 600 + i);
        uint8_t *ab_dst = abuf + (size_t)b * QT_ABUF_BYTES;
        if (!bad) {
          DqPairHalf p0, p1, p2, p3;
          p0.r0 = r[0]; p0.c0 = -r[0] * 8388608.f; p0.r1 = r[1]; p0.c1 = -r[1] * 8388608.f; p0.off2 = of.x;
          p1.r0 = r[2]; p1.c0 = -r[2] * 8388608.f; p1.r1 = r[3]; p1.c1 = -r[3] * 8388608.f; p1.off2 = of.y;
          p2.r0 = r[4]; p2.c0 = -r[4] * 8388608.f; p2.r1 = r[5]; p2.c1 = -r[5] * 8388608.f; p2.off2 = of.z;
          p3.r0 = r[6]; p3.c0 = -r[6] * 8388608.f; p3.r1 = r[7]; p3.c1 = -r[7] * 8388608.f; p3.off2 = of.w;
          uint32_t const src0 = smem_u32(codes) + code_off, dst0 = smem_u32(ab_dst) + dst_off;
#pragma unroll
          for (int ii = 0; ii < 4; ii++) {
            uint32_t m[8];
            if (QB == 8) {
              uint32_t qx, qy;
              asm volatile("ld.shared.v2.u32 {%0,%1}, [%2];" : "=r"(qx), "=r"(qy) : "r"(src0 + ii * 8 * PITCH));
              m[0] = __byte_perm(qx, kDqMagic, 0x7540); m[1] = __byte_perm(qx, kDqMagic, 0x7541);
              m[2] = __byte_perm(qx, kDqMagic, 0x7542); m[3] = __byte_perm(qx, kDqMagic, 0x7543);
              m[4] = __byte_perm(qy, kDqMagic, 0x7540); m[5] = __byte_perm(qy, kDqMagic, 0x7541);
              m[6] = __byte_perm(qy, kDqMagic, 0x7542); m[7] = __byte_perm(qy, kDqMagic, 0x7543);
            } else { // element 2 e in the HIGH nibble of byte e (decompress_kernels.cu:43-50)
              uint32_t q;
              asm volatile("ld.shared.u32 %0, [%1];" : "=r"(q) : "r"(src0 + ii * 8 * PITCH));
#pragma unroll
              for (int e = 0; e < 4; e++) {
                uint32_t const byte = (q >> (8 * e)) & 0xFFu;
                m[2 * e] = (byte >> 4) | kDqMagic;
                m[2 * e + 1] = (byte & 0xFu) | kDqMagic;
              }
            }
#ifdef QT_EXP_NO_DQ // experiment build: the pipeline without the dequantisation arithmetic
            uint32_t const v0 = m[0] ^ m[1] ^ p0.off2, v1 = m[2] ^ m[3] ^ p1.off2, v2 = m[4] ^ m[5], v3 = m[6] ^ m[7];
#else
            uint32_t const v0 = dq_pair(m[0], m[1], p0), v1 = dq_pair(m[2], m[3], p1);
            uint32_t const v2 = dq_pair(m[4], m[5], p2), v3 = dq_pair(m[6], m[7], p3);
#endif
            asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(dst0 + ii * 8 * 128), "r"(v0), "r"(v1), "r"(v2), "r"(v3) : "memory");
          }
        } else {
          qt_slow_convert<QB>(codes, PITCH, reinterpret_cast<__half const *>(meta + meta_off),
                              reinterpret_cast<__half const *>(meta + META_BYTES / 2 + meta_off), ab_dst, g, j, c);
        }
        if (tid == 0) QT_TR(2, 300 + i);
        fence_proxy_async(); // generic-proxy writes -> visible to the tensor core's operand reads
        __syncwarp();
        if (tid == 0) QT_TR(2, 400 + i);
        if (lane == 0) {
          mbar_arrive(&conv_bar[b]);
          if (ss == SUBS - 1) {
            mbar_arrive(&empty_bar[s]);
          }
        }
        } // ss
        if (++s == STAGES) {
          s = 0;
          ph ^= 1u;
        }
      }
      // ----- end of a segment: warps 0-3 read the accumulator (thread = feature row) -----
      if (warp < 4) {
        int const ab = seg & 1;
        if (tid == 0) QT_TR(3, 100 + seg);
        mbar_wait_sleep(&tmem_full_bar[ab], (seg >> 1) & 1);
        if (tid == 0) QT_TR(3, 200 + seg);
        tc_fence_after();
        int const row = warp * 32 + lane, n = w.t * QT_TM + row;
        uint32_t const tsrc = tmem_base + (uint32_t)ab * 64u + ((uint32_t)(warp * 32) << 16);
        bool const head = w.ks0 == 0;
        float *const my = ws + (size_t)cta * (64 * QT_TM) + row; // this CTA's partial slot: [token][128 rows]
        for (int m0 = 0; m0 < NPAD; m0 += 16) {
          uint32_t r[16];
          tmem_ld16(tsrc + (uint32_t)m0, r);
          if (!head) { // park the partial; the CTA that holds the head of this tile adds it
#pragma unroll
            for (int e = 0; e < 16; e++) {
              if (m0 + e < M) {
                __stcg(my + (size_t)(m0 + e) * QT_TM, __uint_as_float(r[e]));
              }
            }
          } else {
            float f[16];
#pragma unroll
            for (int e = 0; e < 16; e++) {
              f[e] = __uint_as_float(r[e]);
            }
            if (w.ks1 < spt) { // the tile runs on into the following CTAs' first segments: add them in CTA order
              int const tile_end = (w.t + 1) * spt;
              int np = 0; // pieces parked by the following CTAs
              while (cta + 1 + np < G && qt_bound(cta + 1 + np, total, G) < tile_end) {
                np++;
              }
              if (m0 == 0) {
                if (tid < np) { // one waiter per piece
                  int seen;
                  do {
                    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(cnt + cta + 1 + tid) : "memory");
                  } while (seen == 0);
                }
                asm volatile("bar.sync 2, 128;" ::: "memory");
              }
              // four pieces' loads in flight at a time (a dependent L2 round trip per piece was a 4-6 us tail)
              for (int p0 = 0; p0 < np; p0 += 4) {
                float pv[4][16];
#pragma unroll
                for (int q = 0; q < 4; q++) {
                  float const *theirs = ws + (size_t)(cta + 1 + min(p0 + q, np - 1)) * (64 * QT_TM) + row;
#pragma unroll
                  for (int e = 0; e < 16; e++) {
                    pv[q][e] = (p0 + q < np && m0 + e < M) ? __ldcg(theirs + (size_t)(m0 + e) * QT_TM) : 0.f;
                  }
                }
#pragma unroll
                for (int q = 0; q < 4; q++) {
                  if (p0 + q < np) {
#pragma unroll
                    for (int e = 0; e < 16; e++) {
                      f[e] += pv[q][e];
                    }
                  }
                }
              }
            }
            float const bv = bias ? __half2float(bias[n]) : 0.f;
#pragma unroll
            for (int e = 0; e < 16; e++) {
              if (m0 + e < M) {
                out[(size_t)(m0 + e) * N + n] = __float2half_rn(f[e] + bv);
              }
            }
          }
        }
        tc_fence_before();
        if (!head) {
          __threadfence();
          asm volatile("bar.sync 2, 128;" ::: "memory");
          if (tid == 0) {
            atomicExch(cnt + cta, 1);
          }
        } else if (w.ks1 < spt) { // flags back to zero for the next launch (their writers raise them once per launch)
          asm volatile("bar.sync 2, 128;" ::: "memory");
          int const tile_end = (w.t + 1) * spt;
          int const cq = cta + 1 + tid;
          if (cq < G && qt_bound(cq, total, G) < tile_end) {
            cnt[cq] = 0;
          }
        }
        __syncwarp();
        if (tid == 0) QT_TR(3, 300 + seg);
        if (lane == 0) {
          mbar_arrive(&tmem_empty_bar[ab]);
        }
      }
      seg++;
    }
  }
  tc_fence_before();
#ifdef FFB_QT_TRACE_BUILD
  if (tid == 0) {
    long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    g_qt_cta[blockIdx.x][1] = t_;
  }
#endif
  __syncthreads();
  if (warp == QT_MMA_WARP) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128u));
  }
#ifdef FFB_QT_TRACE_BUILD
  if (tid == 0) {
    long long t_;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));
    g_qt_cta[blockIdx.x][2] = t_;
  }
#endif
}

} // namespace

// one partial slot (64 tokens x 128 floats) and one flag per CTA
size_t linear_qtc_ws_bytes() { return (size_t)kNumSMs * 64 * QT_TM * sizeof(float) + (size_t)kNumSMs * sizeof(int); }

// half activations, int8 / int4 blobs, up to 64 tokens.  OFF unless FFB_QTC=1 (FFB_LINEAR=qtc in the tests): the kernel is
// bit-exact (tests/test_gpu_linear_qtc.py) but measured SLOWER than the mma.sync kernels at the 7B decode shapes - 27.5
// against 23.5 us (qkv_proj, int8, 16 tokens), 18.4 against 11.3 us (o_proj) - and its time does not move with the
// dequantisation arithmetic (removed: 26.1 us), the number of MMAs per k-step (1 instead of 8: 25.4 us), the operand
// buffers (2 / 4), the bytes per row visit (128 / 256), the number of copy instructions per stage (11 / 4) or L2
// residency of the blob: ~1 us per 128-wide k-step per SM plus a 6-8 us tail in the CTAs that merge the cut tiles
// (per-CTA start / end stamps, -DFFB_QT_TRACE_BUILD).  The single-thread producer / MMA-issuer roles next to sixteen
// busy converter warps are the suspects; it stays in the tree as the starting point for that work.
int linear_qtc_supported(int dtype, int quant, int M, int K, int N) {
  char const *e = env("FFB_QTC");
  char const *f = env("FFB_LINEAR");
  if (!((e && e[0] == '1') || (f && strcmp(f, "qtc") == 0))) {
    return 0;
  }
  return dtype == FFB_DT_HALF && (quant == FFB_DT_INT8 || quant == FFB_DT_INT4) && M >= 1 && M <= 64 && K % QT_KS == 0 &&
         N % QT_TM == 0 && (long)(N / QT_TM) * (K / QT_KS) >= kNumSMs;
}

int launch_linear_qtc(int quant, void const *in, void *out, void const *blob, void const *bias, int M, int K, int N,
                      void *ws, cudaStream_t s) {
  int const npad = (M + 15) / 16 * 16;
  int const bits = quant == FFB_DT_INT8 ? 8 : 4;
  int rc;
  // k-steps per stage: a stage's code box spans SUBS * 128 k of every row; the TMA box is at most 256 bytes wide
  // (int8: 2 steps, int4: 4) and three stages must fit beside the two operand buffers
  int subs = 2;
  if (char const *e = env("FFB_QTC_SUBS")) {
    int const v = atoi(e);
    if (v == 1 || v == 2) {
      subs = v;
    }
  }
  size_t const fixed = (size_t)QT_NBUF * QT_ABUF_BYTES + 1024;
  auto stage_bytes = [&](int sb) { return (size_t)sb * ((size_t)npad * 256 + (size_t)QT_TM * QT_KS * bits / 8 + QT_META_BYTES); };
  while (subs > 1 && ((K / QT_KS) % subs != 0 || 2 * stage_bytes(subs) + fixed > ((size_t)212 << 10))) {
    subs /= 2;
  }
  size_t const stage = stage_bytes(subs);
  int const stages = (int)std::min<size_t>(QT_MAX_STAGES, (((size_t)212 << 10) - fixed) / stage);
  size_t const smem = (size_t)stages * stage + fixed;
  float *wsf = (float *)ws;
  int *cnt = (int *)((char *)ws + (size_t)kNumSMs * 64 * QT_TM * sizeof(float));
  int const total = (N / QT_TM) * (K / (QT_KS * subs));
  int const grid = std::min(kNumSMs, total);
  CUtensorMap tmW;
  rc = make_tensor_map_u8(&tmW, blob, N, (long)K * bits / 8, QT_TM, subs * QT_KS * bits / 8);
  if (rc != FFB_OK) {
    return rc;
  }
  // activations [M][K] viewed as [K / 64 atoms][M rows][64]: one box = 2 SUBS atoms of npad rows (rows >= M read zero)
  CUtensorMap tmX, tmM;
  {
    unsigned long long dims[3] = {64ull, (unsigned long long)M, (unsigned long long)(K / 64)};
    unsigned long long strides[2] = {(unsigned long long)K * 2, 128ull};
    unsigned box[3] = {64u, (unsigned)npad, (unsigned)(2 * subs)};
    rc = make_tensor_map_nd(&tmX, 2, true, 3, in, dims, strides, box, true);
    if (rc != FFB_OK) {
      return rc;
    }
    // offsets then scales: [2 N / 32 rows][K] 16-bit words right behind the codes
    size_t const nvalues = (size_t)N * K;
    void const *meta = (char const *)blob + (bits == 4 ? nvalues / 2 : nvalues);
    unsigned long long mdims[2] = {(unsigned long long)K, (unsigned long long)(2 * (N / 32))};
    unsigned long long mstrides[1] = {(unsigned long long)K * 2};
    unsigned mbox[2] = {(unsigned)(subs * QT_KS), 4u};
    rc = make_tensor_map_nd(&tmM, 2, false, 2, meta, mdims, mstrides, mbox, false);
    if (rc != FFB_OK) {
      return rc;
    }
  }
#define FFB_QTC(QBv, SUBSv)                                                                                          \
  do {                                                                                                               \
    FFB_SET_MAX_SMEM_ONCE((linear_qtc_kernel<QBv, SUBSv>), ((size_t)213 << 10));                                     \
    linear_qtc_kernel<QBv, SUBSv><<<grid, QT_THREADS, smem, s>>>(tmW, tmX, tmM, (uint8_t const *)blob, (__half const *)bias, \
                                                                 (__half *)out, wsf, cnt, M, N, K, npad, stages);   \
  } while (0)
  if (bits == 8) {
    if (subs == 2) FFB_QTC(8, 2); else FFB_QTC(8, 1);
  } else {
    if (subs == 2) FFB_QTC(4, 2); else FFB_QTC(4, 1);
  }
#undef FFB_QTC
  FFB_LAUNCH_CHECK();
#ifdef FFB_QT_TRACE_BUILD
  static int calls = 0;
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing(s, &cap);
  if (cap == cudaStreamCaptureStatusNone && ++calls % 8 == 0) {
    long long h[4][64][2];
    int n[4];
    cudaStreamSynchronize(s);
    cudaMemcpyFromSymbol(h, g_qt_trace, sizeof(h));
    cudaMemcpyFromSymbol(n, g_qt_n, sizeof(n));
    long long hc[160][3];
    cudaMemcpyFromSymbol(hc, g_qt_cta, sizeof(hc));
    long long t0 = hc[0][0];
    for (int c = 0; c < grid; c++) {
      t0 = std::min(t0, hc[c][0]);
    }
    fprintf(stderr, "QTCTA start/warp0-done/end (ns from first start):");
    for (int c = 0; c < grid; c += 1) {
      fprintf(stderr, " %d:%lld/%lld/%lld", c, hc[c][0] - t0, hc[c][1] - t0, hc[c][2] - t0);
    }
    fprintf(stderr, "\n");
    for (int r = 0; r < 4; r++) {
      fprintf(stderr, "QTTRACE role %d:", r);
      for (int e = 0; e < n[r] && e < 64; e++) {
        fprintf(stderr, " %lld@%lld", h[r][e][0], h[r][e][1] - h[0][0][1]);
      }
      fprintf(stderr, "\n");
    }
  }
  int zero[4] = {0, 0, 0, 0};
  cudaMemcpyToSymbolAsync(g_qt_n, zero, sizeof(zero), 0, cudaMemcpyHostToDevice, s);
#endif
  return FFB_OK;
}

} // namespace ffb
