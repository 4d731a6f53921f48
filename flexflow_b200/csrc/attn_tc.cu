// attn_tc.cu — tcgen05 / TMA multi-query attention over a shared KV stream: tree-verify (bit mask), beam / spec
// (bit mask with the SSM's query-bit rule), chunked-prefill prompt (causal) and GQA decode.
// Replaces compute_attention_kernel_fused_kernel (src/ops/tree_inc_multihead_self_attention.cu:35-333),
// compute_spec_inc_attention_kernel_generation_kernel (src/ops/spec_inc_multihead_self_attention.cu:36-309)
// and the cuBLAS + cuDNN prompt path (src/ops/inc_multihead_self_attention.cu:98-366).
//
// PERSISTENT, one CTA per SM.  A work unit is a 128-row query tile (query tokens x the q heads sharing one kv
// head) of one request; its K/V is cut into 128-key tiles.  The host flattens all (unit, key tile) pairs of the
// batch and gives every CTA one contiguous range of that list (stream-K), so the SMs finish together whatever
// the request lengths are; a unit that straddles CTAs is merged through split records (exact log-sum-exp merge
// by the last CTA to finish the row, combine_record).  Every K/V tile is read from HBM exactly once:
//   warp 8 (one lane)  TMA producer: K and V tiles straight from the [request][kv_head][cache_len][Dh] cache into
//                      a 192 KB ring of 128B-swizzled buffers (K0 V0 K1 V1 ...); it runs ahead across unit
//                      boundaries, so the HBM stream never drains while a unit's epilogue / the next Q load runs
//   warp 9 (one lane)  MMA issuer: S = Q K^T (UMMA 128x128x16, Q from TENSOR MEMORY, K from smem K-major) into one
//                      of two TMEM score buffers, then O += P V (P from tensor memory - it overwrites the first
//                      64 columns of its own score buffer - and V consumed MN-major exactly as TMA laid it down)
//   warps 0-15         FOUR threads per query row (warps w, w+4, w+8, w+12 share TMEM lane quarter w % 4; each takes
//                      32 of the 128 keys and a quarter of the output columns): tcgen05.ld the scores, scale + kv_len /
//                      causal / tree-bit mask, online softmax in the log2 domain against a LAZY reference maximum (the
//                      TMEM accumulator is rescaled only when a row maximum grows by more than 2^8, so P stays <= 256
//                      and most tiles need no correction), P back into TMEM (tcgen05.st).  A 64-token MHA tree fills
//                      only 64 of a unit's 128 rows, i.e. lane quarters 0 and 1: with two threads per row four of
//                      eight warps did all the work at 64 elements each and the tile loop ran at 1.2 us per tile,
//                      bound by their issue rate and exposed tcgen05.ld / st latency (phase trace, profiles/); with
//                      four threads per row eight warps share it at 32 elements each.
// The tree mask of a query row arrives TRANSPOSED in its QueryDesc (one word per row, written by the prep kernel).
// Shared memory holds nothing but the K/V stream.  HBM-bound: 64 KB of K/V per tile per SM
// (~2900 cycles at the HBM rate) against ~1000 tensor-pipe and ~1000 MUFU cycles.
#include <algorithm>
#include <cstdio>
#include <type_traits>

#include "attn_state.h"
#include "tc_common.cuh"

namespace ffb {

using namespace tc;

namespace {

constexpr int TBM = 128, TBN = 128;
constexpr int ATC_NQ = 4;                               // threads per query row = column quarters of a key tile
constexpr int ATC_SOFTMAX_THREADS = TBM * ATC_NQ;       // warps 0 .. 4 NQ - 1
constexpr int ATC_PRODUCER_WARP = 4 * ATC_NQ, ATC_MMA_WARP = 4 * ATC_NQ + 1;
constexpr int ATC_THREADS = ATC_SOFTMAX_THREADS + 64;
constexpr float kRescaleThreshold = 8.f; // log2 domain
constexpr int ATC_MAX_ENT = 128;

// One entry per (request, 128-row query tile); it stands for num_kv_heads units of `ntiles` key tiles each.
// Flattened tile index of (entry e, kv head h, tile t) = cum[e] + h * ntiles[e] + t.
struct AtcEntry {
  int cum;        // flattened index of this entry's first tile
  int first_tok;  // first_token_offset_in_batch of the request
  int ntcs;       // causalMask[slot].non_tree_cache_size
  short slot, mt; // request slot, row-tile index inside the request
  short ntiles;   // key tiles of 128 covering the longest row of the tile
  short qlen;     // query tokens of the request in this batch
  int first_new;  // first cache row of the request that THIS batch's prep kernel writes (rows below it are older)
};
struct AtcSched {
  int n_ent;     // entries in use
  int total;     // total flattened tiles
  int rg;        // tiles per CTA (CTA c owns [c*rg, min(total, (c+1)*rg)))
  int maxpieces; // stride of the split records of one (token, head)
  AtcEntry ent[ATC_MAX_ENT];
};

template <int DH> struct AtcCfg {
  static constexpr int KH = DH / 64;                    // 64-element (128-byte) column blocks per row
  static constexpr uint32_t HALF_BYTES = 128 * 128;     // one [128 keys][64 cols] swizzled block
  static constexpr uint32_t KV_BYTES = KH * HALF_BYTES; // one K tile or one V tile
  static constexpr int NBUF = (160 * 1024) / KV_BYTES;  // ring slots; K and V tiles alternate: K0 V0 K1 V1 ...
  // next segment's Q rows, staged by cp.async as [softmax warp][lane][DH / 2 bytes]; the same per-warp blocks serve as
  // the output staging tiles of the epilogue (Q has been moved to tensor memory by then)
  static constexpr uint32_t Q_BYTES = 128 * DH * 2;
  static constexpr size_t SMEM = (size_t)NBUF * KV_BYTES + Q_BYTES + 1024; // ring + Q / output staging + alignment
};

// x / d for 0 <= x < 2^22, 0 < d, rd = 1.f / d: a float estimate and a +-1 fix-up (8 instructions; an integer division by a
// run-time value is ~25, and this code runs at every unit boundary in all 18 warps, out of a cold instruction cache)
__device__ __forceinline__ int fast_div(int x, int d, float rd) {
  int q = (int)((float)x * rd);
  int const r = x - q * d;
  q += r < 0 ? -1 : (r >= d ? 1 : 0);
  return q;
}

// Walks the segments of one CTA: maximal runs of consecutive flattened tiles inside one unit.  The position inside the
// schedule (entry, kv head, tile) is kept incrementally: a segment ends at the end of its unit or of the CTA's range.
struct SegWalk {
  int x, x_end; // flattened position, end of this CTA's range
  int e;        // entry cursor
  int nkv;
  int kvh_next, t_next; // position of x inside entry e
  // current segment
  int kvh, t0, t1;
  __device__ __forceinline__ void init(AtcSched const &sc, int nkv_) {
    x = (int)blockIdx.x * sc.rg;
    x_end = min(sc.total, x + sc.rg);
    nkv = nkv_;
    e = 0;
    while (e + 1 < sc.n_ent && sc.ent[e + 1].cum <= x) {
      e++;
    }
    int const nt = sc.ent[e].ntiles;
    int const r = x - sc.ent[e].cum;
    kvh_next = fast_div(r, nt, __frcp_rn((float)nt));
    t_next = r - kvh_next * nt;
  }
  __device__ __forceinline__ bool next(AtcSched const &sc) {
    if (x >= x_end) {
      return false;
    }
    int nt = sc.ent[e].ntiles;
    if (t_next == nt) { // the previous segment ran to the end of its unit
      t_next = 0;
      if (++kvh_next == nkv) {
        kvh_next = 0;
        e++;
        nt = sc.ent[e].ntiles;
      }
    }
    kvh = kvh_next;
    t0 = t_next;
    t1 = min(nt, t0 + (x_end - x));
    x += t1 - t0;
    t_next = t1;
    return true;
  }
};

template <typename T, int DH>
// 18 warps are allocated as 20 (warp allocation granularity 4): 65536 / 640 = 102 registers per thread at most
__global__ void __launch_bounds__(ATC_THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
               const __grid_constant__ AtcSched sched_param, T const *__restrict__ q_buf,
               QueryDesc const *__restrict__ qdesc, ffb_combined_batch_config_meta const *__restrict__ meta,
               float *__restrict__ partial, int *__restrict__ split_counter, T *__restrict__ out, int nq, int nkv,
               int cache_len, float scale_log2, int kv_evict_first, int q_first, long long *trace) {
  using Cfg = AtcCfg<DH>;
  constexpr int KH = Cfg::KH, NBUF = Cfg::NBUF;
  constexpr uint32_t HALF = Cfg::HALF_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t kv_full[NBUF], kv_empty[NBUF];
  __shared__ __align__(8) uint64_t s_full[2], o_full, p_full, q_full, q_issued;
  __shared__ float s_xch[2][ATC_NQ][TBM]; // [tile parity][column quarter][row]: partial row maxima / final row sums
  __shared__ uint32_t tmem_base_smem;
  __shared__ int s_rowlast[TBM];
  __shared__ AtcSched s_sched;

  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const G = nq / nkv;
#ifdef FFB_ATC_TRACE_BUILD
  // nvcc -DFFB_ATC_TRACE_BUILD + env FFB_ATC_TRACE=<n>: CTA 0 records (event id, globaltimer) pairs of its n-th launch;
  // role r (0 producer, 1 MMA issuer, 2 softmax thread 0) writes into trace[r*512 ...]
  int tr_n = 0;
  auto tr = [&](int role, int ev) {
    if (trace && blockIdx.x == 0 && tr_n < 255) {
      long long t;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
      trace[role * 512 + 2 * tr_n] = ev;
      trace[role * 512 + 2 * tr_n + 1] = t;
      tr_n++;
    }
  };
#define FFB_TR(role, ev) tr(role, ev)
  if (trace && tid == 0) { // every CTA: start / end stamps and the SM it ran on, in trace[1536 + 3 * cta ...]
    long long t;
    unsigned smid;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    trace[1536 + 3 * blockIdx.x] = t;
    trace[1536 + 3 * blockIdx.x + 2] = smid;
  }
#else
#define FFB_TR(role, ev) ((void)0)
#endif
  // 128B-swizzled operand blocks need 1024-byte alignment
  uint8_t *sKV = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *const sQ = sKV + (size_t)NBUF * Cfg::KV_BYTES;     // the NEXT segment's Q / this segment's output staging

  // The tile schedule is walked by all three roles at every unit boundary.  In parameter (constant) space each
  // first touch of a line is a ~0.7 us miss and the walks are dependent chains (2 us per boundary, measured): copy
  // the entries in use to shared memory once.
  {
    int const *src = reinterpret_cast<int const *>(&sched_param);
    int *dst = reinterpret_cast<int *>(&s_sched);
    int const words = 4 + sched_param.n_ent * (int)(sizeof(AtcEntry) / 4);
    for (int i = tid; i < words; i += ATC_THREADS) {
      dst[i] = src[i];
    }
  }
  __syncthreads();
  AtcSched const &sched = s_sched;
  if (tid == 0) {
    // every CTA of this (single-wave) grid is resident: the next kernel of the stream - the o_proj GEMM, launched as a
    // programmatic dependent - may start its CTAs on SMs as they free up and request its first weight tiles
    pdl_trigger();
  }

  if (warp == ATC_PRODUCER_WARP) {
    // ===== TMA producer =====  (never waits for the rest of the CTA: the stream starts at once)
    if (lane == 0) {
      for (int sb = 0; sb < NBUF; sb++) {
        mbar_init(&kv_full[sb], 1);
        mbar_init(&kv_empty[sb], 1);
      }
      mbar_init(&s_full[0], 1);
      mbar_init(&s_full[1], 1);
      mbar_init(&o_full, 1);
      mbar_init(&p_full, ATC_SOFTMAX_THREADS);
      mbar_init(&q_full, ATC_SOFTMAX_THREADS);
      mbar_init(&q_issued, ATC_SOFTMAX_THREADS / 32);
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    }
    __syncwarp();
    asm volatile("barrier.cta.arrive 0, %0;" ::"n"(ATC_THREADS) : "memory");
    if (lane == 0) {
      SegWalk w;
      w.init(sched, nkv);
      // ring cursor: slot and phase kept incrementally (j % NBUF, j / NBUF would make the compiler unroll the loop
      // NBUF-fold: 8 KB of a 32 KB instruction cache for one lane's loop)
      int sb = 0;
      uint32_t ph = 1;
      // K / V rows are read once: evict-first in L2, so that the stream does not push out what is re-read (code,
      // descriptors, Q, split records)
      uint64_t policy = 0;
      if (kv_evict_first) {
        asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
      }
      // Programmatic dependent launch: this grid may start while the prep kernel (RoPE + KV append of THIS batch) is
      // still running.  Cache rows below first_new are older than that kernel, so their tiles stream at once; the
      // first tile that reaches into the new rows waits for the prep grid (griddepcontrol.wait, once).
      bool prep_done = false;
      int j = 0;
      while (w.next(sched)) {
        int const row_base = ((int)sched.ent[w.e].slot * nkv + w.kvh) * cache_len;
        int const first_new = sched.ent[w.e].first_new;
#pragma unroll 1
        for (int t = w.t0; t < w.t1; t++) {
          if (!prep_done && (t + 1) * TBN > first_new) {
            asm volatile("griddepcontrol.wait;" ::: "memory");
            prep_done = true;
          }
#pragma unroll 1
          for (int kv = 0; kv < 2; kv++, j++) {
            mbar_wait(&kv_empty[sb], ph);
            FFB_TR(0, j);
            mbar_expect_tx(&kv_full[sb], Cfg::KV_BYTES);
            CUtensorMap const *tm = kv ? &tmV : &tmK;
#pragma unroll
            for (int hh = 0; hh < KH; hh++) {
              if (kv_evict_first) {
                tma_load_2d_hint(sKV + sb * Cfg::KV_BYTES + hh * HALF, tm, hh * 64, row_base + t * TBN, &kv_full[sb], policy);
              } else {
                tma_load_2d(sKV + sb * Cfg::KV_BYTES + hh * HALF, tm, hh * 64, row_base + t * TBN, &kv_full[sb]);
              }
            }
            if (++sb == NBUF) {
              sb = 0;
              ph ^= 1u;
            }
            if (j == 0 && q_first) {
              // The first Q rows are a few KB on the critical path (no QK^T without them); queued behind this CTA's
              // 160 KB of K / V they arrive 2-3 us later.  One tile goes out, then the softmax warps' loads.
              mbar_wait(&q_issued, 0);
            }
          }
        }
      }
    }
  } else {
    if (warp == ATC_MMA_WARP) {
      // TMEM: score buffers S0 S1 (P overwrites the first 64 columns of its S), output accumulator, Q operand
      asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
      asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    asm volatile("barrier.cta.sync 0, %0;" ::"n"(ATC_THREADS) : "memory");
    tc_fence_after();
    uint32_t const tmem_base = tmem_base_smem;
    uint32_t const tmem_S = tmem_base, tmem_O = tmem_base + 256, tmem_Q = tmem_base + 384;

    if (warp == ATC_MMA_WARP) {
      if (lane == 0) {
        // ===== MMA issuer =====
        // cute::UMMA::InstrDescriptor: c_format F32 (1) @4, a/b format @7/@10 (F16 0, BF16 1), b_major @16
        // (1 = MN-major), N>>3 @17, M>>4 @24.  The A operand (Q, then P) is in tensor memory: 8 columns per K=16 step.
        constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
        constexpr uint32_t idesc_qk = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TBN >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);
        constexpr uint32_t idesc_pv = (1u << 4) | (fmt << 7) | (fmt << 10) | (1u << 16) | ((uint32_t)(DH >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);
        auto issue_qk = [&](int g) { // g: CTA-wide tile counter
          int const j = 2 * g, sb = j % NBUF;
          mbar_wait(&kv_full[sb], (j / NBUF) & 1);
          FFB_TR(1, 1000 + g);
          tc_fence_after();
          uint32_t const kb = smem_u32(sKV + sb * Cfg::KV_BYTES);
#pragma unroll
          for (int k = 0; k < DH / 16; k++) {
            uint64_t const db = umma_desc_sw128(kb + (k >> 2) * HALF) + (uint64_t)((k & 3) * 2);
            tc_mma_f16_ts(tmem_S + (uint32_t)(g & 1) * 128u, tmem_Q + (uint32_t)(k * 8), db, idesc_qk, k != 0);
          }
          tc_commit(&kv_empty[sb]);
          tc_commit(&s_full[g & 1]);
        };
        SegWalk w;
        w.init(sched, nkv);
        int g0 = 0, seg = 0;
        while (w.next(sched)) {
          int const ntiles = w.t1 - w.t0;
          if (g0 > 0) {
            mbar_wait(&o_full, (g0 - 1) & 1); // every MMA of the previous segment has completed
          }
          FFB_TR(1, 4000 + seg);
          mbar_wait(&q_full, seg & 1);        // Q of this segment is in tensor memory, O of the previous one was read out
          tc_fence_after();
          // one loop body for the whole segment (it = -2, -1: the two look-ahead QK^T), so that issue_qk and the PV
          // block are emitted once: the issuer's code stays small in the instruction cache
#pragma unroll 1
          for (int it = -2; it < ntiles; it++) {
            int const g = g0 + it;
            if (it >= 0) {
              int const j = 2 * g + 1, sb = j % NBUF;
              mbar_wait(&p_full, g & 1);
              FFB_TR(1, 2000 + g);
              mbar_wait(&kv_full[sb], (j / NBUF) & 1);
              FFB_TR(1, 3000 + g);
              tc_fence_after();
              uint32_t const vb = smem_u32(sKV + sb * Cfg::KV_BYTES);
              uint32_t const tP = tmem_S + (uint32_t)(g & 1) * 128u;
#pragma unroll
              for (int k = 0; k < TBN / 16; k++) {
                uint64_t const db = umma_desc_mn_sw128(vb + k * 2048, HALF, 1024);
                tc_mma_f16_ts(tmem_O, tP + (uint32_t)(k * 8), db, idesc_pv, (it | k) != 0);
              }
              tc_commit(&kv_empty[sb]);
              tc_commit(&o_full);
            }
            if (it + 2 < ntiles) {
              if (it >= 0) {
                mbar_wait(&o_full, g & 1); // P of this tile has been consumed: its score buffer may be overwritten
              }
              issue_qk(g + 2);
            }
          }
          g0 += ntiles;
          seg++;
        }
      }
    } else {
      // ===== softmax warps: thread (row, hc) owns keys [32 hc, 32 hc + 32) of every tile and DH / 4 output columns =====
      int const row = tid & (TBM - 1), hc = tid >> 7;
      uint32_t const lane_sel = (uint32_t)((warp & 3) * 32) << 16;
      uint32_t const quad_bar = 1 + (warp & 3); // named barrier shared by the NQ warps of one lane quarter
      constexpr int KC = TBN / ATC_NQ;          // keys (score columns) per thread: 32
      constexpr int PC = KC / 2;                // packed P columns per thread: 16
      constexpr int OC = DH / ATC_NQ;           // output columns per thread: 32 (Dh 128) / 16 (Dh 64)
      constexpr int QW = DH / 2 / ATC_NQ;       // 32-bit words of Q per thread: 16 / 8
      constexpr int QCH = QW / 4;               // 16-byte chunks of this thread's Q / output slice: 4 / 2
      static_assert(KC == 32 && (OC == 32 || OC == 16), "tile shape");
      uint32_t const tO = tmem_O + lane_sel + (uint32_t)(hc * OC);
      uint32_t const tq = tmem_Q + lane_sel + (uint32_t)(hc * QW);
      // this warp's 32 x (DH / 2)-byte block of the Q / staging area; 16-byte chunks XOR-swizzled by lane so that both the
      // thread-per-row accesses and the 4 (2)-lanes-per-row read-back of the epilogue are conflict-free
      uint8_t *const wblk = sQ + (size_t)warp * (32 * QCH * 16);
      int const swz = QCH == 4 ? ((lane >> 1) & 3) : ((lane >> 2) & 1);

      // Row data of a segment: fetched one segment ahead (global loads in flight during the tile loop), so a unit
      // boundary costs no exposed global-memory latency.
      int n_th = -1, n_kvl = 0, n_bit = -1;
      unsigned long long n_vis = 0;
      float const rcp_G = __frcp_rn((float)G), rcp_rg = __frcp_rn((float)sched.rg);
      auto fetch_rows = [&](SegWalk const &sw) {
        AtcEntry const &en = sched.ent[sw.e];
        int const R = en.mt * TBM + row;
        n_th = -1;
        n_kvl = 0;
        n_bit = -1;
        n_vis = 0;
        if (R < en.qlen * G) {
          int const ti = fast_div(R, G, rcp_G);
          int const tok = en.first_tok + ti;
          n_th = tok * nq + sw.kvh * G + (R - ti * G);
          QueryDesc const qd = qdesc[tok];
          n_kvl = qd.kv_len;
          n_bit = qd.bit;
          n_vis = qd.vis;
          // this thread's quarter row of Q -> shared memory, asynchronously (no registers held while the current
          // segment streams; under load an L2 round trip at the unit boundary costs 2-3 us)
          uint32_t const dst = smem_u32(wblk) + (uint32_t)lane * (QCH * 16);
          T const *src = q_buf + (size_t)n_th * DH + hc * (DH / ATC_NQ);
#pragma unroll
          for (int c = 0; c < QCH; c++) {
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + (uint32_t)((c ^ swz) << 4)), "l"(src + c * 8) : "memory");
          }
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
      };
      // Q -> tensor memory (all QK^T MMAs of the previous segment have completed: their s_full phases were observed)
      auto publish_rows = [&]() {
        uint32_t n_q[QW];
        asm volatile("cp.async.wait_group 0;" ::: "memory");
        uint8_t const *src = wblk + (size_t)lane * (QCH * 16);
#pragma unroll
        for (int c = 0; c < QCH; c++) {
          uint4 const v = n_th >= 0 ? *reinterpret_cast<uint4 const *>(src + ((c ^ swz) << 4)) : make_uint4(0u, 0u, 0u, 0u);
          n_q[c * 4 + 0] = v.x; n_q[c * 4 + 1] = v.y; n_q[c * 4 + 2] = v.z; n_q[c * 4 + 3] = v.w;
        }
        tmem_stN<QW>(tq, n_q);
        tmem_st_wait();
        // (release) the first QK^T of the segment is issued after all arrivals: no CTA-wide barrier at the unit boundary
        tc_fence_before();
        mbar_arrive(&q_full);
      };

      // q_buf / qdesc come from the prep kernel of this batch (see the producer)
      asm volatile("griddepcontrol.wait;" ::: "memory");
      SegWalk w;
      w.init(sched, nkv);
      bool have = w.next(sched);
      int g0 = 0;
      if (tid == 0) FFB_TR(2, 100);
      if (have) {
        fetch_rows(w);
      }
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&q_issued); // the producer holds its stream back until the first Q loads are queued ahead of it
      }
      if (have) {
        if (tid == 0) FFB_TR(2, 101);
        publish_rows();
      }
      if (tid == 0) FFB_TR(2, 102);
      while (have) {
        AtcEntry const &en = sched.ent[w.e];
        int const ntiles = w.t1 - w.t0, ntcs = en.ntcs, t_first = w.t0;
        // the pieces of this unit: CTAs whose ranges intersect its flattened tiles
        int const U0 = en.cum + w.kvh * en.ntiles;
        int const c_first = fast_div(U0, sched.rg, rcp_rg);
        int const npieces = fast_div(U0 + en.ntiles - 1, sched.rg, rcp_rg) - c_first + 1;
        int const piece = (int)blockIdx.x - c_first;
        int const th = n_th, kvl = n_kvl, bit = n_bit;
        // tree / beam rows: bit i of row_vis says whether tree position i (key ntcs + i) is visible to this row
        // (mask rule tree_…cu:165-169 / spec_…cu:165-167 with 1ULL semantics), transposed by the prep kernel
        unsigned long long const row_vis = n_vis;
        bool const warp_active = __any_sync(0xffffffffu, th >= 0); // padding-only warps just keep the barriers moving
        if (tid == 0) FFB_TR(2, 105);
        have = w.next(sched);
        if (tid == 0) FFB_TR(2, 106 + (have ? 0 : 1));
        if (have) {
          fetch_rows(w); // next segment's rows: loads stay in flight while this segment streams
        }
        if (tid == 0) FFB_TR(2, 104);
        float m_ref = -INFINITY, l = 0.f;

        for (int it = 0; it < ntiles; it++) {
          int const g = g0 + it, b = g & 1;
          int const kbase = (t_first + it) * TBN + hc * KC; // first key of this thread's columns
          uint32_t const tS = tmem_S + (uint32_t)b * 128u + lane_sel;
          bool const need_mask = __any_sync(0xffffffffu, th >= 0 && (kbase + KC > kvl || (bit >= 0 && kbase + KC > ntcs)));
          float sc[KC];
          if (tid == 0) FFB_TR(2, 1000 + g);
          mbar_wait(&s_full[b], (g >> 1) & 1);
          if (tid == 0) FFB_TR(2, 2000 + g);
          tc_fence_after();
          float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
          if (warp_active) {
            uint32_t r0[32];
            tmem_ld32(tS + (uint32_t)(hc * KC), r0);
#pragma unroll
            for (int e = 0; e < 32; e++) { // raw scores; the (positive) scale is folded into the exponent below
              sc[e] = __uint_as_float(r0[e]);
            }
            if (need_mask) {
              // visibility of this thread's 32 keys as one word: kv_len / causal bound, then the tree bits
              int const lim = kvl - kbase;
              uint32_t V = lim >= 32 ? ~0u : (lim <= 0 ? 0u : ((1u << lim) - 1u));
              if (bit >= 0) {
                int const off = kbase - ntcs; // tree position of column 0
                uint32_t tv;
                if (off >= 64) {
                  tv = 0u;
                } else if (off >= 0) {
                  tv = (uint32_t)(row_vis >> off);
                } else if (off > -32) {
                  tv = (uint32_t)((row_vis << (-off)) | ((1ULL << (-off)) - 1ULL));
                } else {
                  tv = ~0u;
                }
                V &= tv;
              }
#pragma unroll
              for (int e = 0; e < 32; e++) {
                sc[e] = (V >> e) & 1u ? sc[e] : -INFINITY;
              }
            }
#pragma unroll
            for (int e = 0; e < KC; e += 4) {
              mx0 = fmaxf(mx0, sc[e]);
              mx1 = fmaxf(mx1, sc[e + 1]);
              mx2 = fmaxf(mx2, sc[e + 2]);
              mx3 = fmaxf(mx3, sc[e + 3]);
            }
          }
          float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3)) * scale_log2; // log2 domain from here on
          s_xch[b][hc][row] = mx;
          asm volatile("bar.sync %0, %1;" ::"r"(quad_bar), "n"(32 * ATC_NQ) : "memory");
#pragma unroll
          for (int o = 1; o < ATC_NQ; o++) {
            mx = fmaxf(mx, s_xch[b][(hc + o) & (ATC_NQ - 1)][row]);
          }
          // lazy reference maximum: move it (and rescale the accumulator) only when the row maximum outgrows it by 2^8
          bool const grow = mx > m_ref + kRescaleThreshold || (m_ref == -INFINITY && mx > -INFINITY);
          float const alpha = grow ? (m_ref == -INFINITY ? 1.f : ffb_ex2(m_ref - mx)) : 1.f;
          if (grow) {
            m_ref = mx;
          }
          float const base = m_ref == -INFINITY ? 0.f : m_ref;
          if (it > 0 && __any_sync(0xffffffffu, alpha != 1.f)) {
            // rescale the accumulator rows whose reference moved; needs the PV MMA of the previous tile to be complete
            // (the MMA of this tile cannot start before this thread arrives on p_full)
            mbar_wait(&o_full, (g - 1) & 1);
            tc_fence_after();
            uint32_t r[OC];
            tmem_ldN<OC>(tO, r);
#pragma unroll
            for (int e = 0; e < OC; e++) {
              r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
            }
            tmem_stN<OC>(tO, r);
            tmem_st_wait();
          }
          if (warp_active) {
            // P (16-bit pairs) overwrites the first 64 columns of this tile's score buffer: this thread's 32 keys are
            // columns [16 hc, 16 hc + 16); every thread of the row finished reading its scores before the barrier above
            float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
            uint32_t pk[PC];
#pragma unroll
            for (int e = 0; e < KC; e += 4) {
              float f0 = fmaf(sc[e], scale_log2, -base), f1 = fmaf(sc[e + 1], scale_log2, -base);
              float f2 = fmaf(sc[e + 2], scale_log2, -base), f3 = fmaf(sc[e + 3], scale_log2, -base);
              f0 = ffb_ex2(f0);
              f1 = ffb_ex2(f1);
              f2 = ffb_ex2(f2);
              f3 = ffb_ex2(f3);
              s0 += f0;
              s1 += f1;
              s2 += f2;
              s3 += f3;
              pk[e / 2] = DT<T>::pack(f0, f1);
              pk[e / 2 + 1] = DT<T>::pack(f2, f3);
            }
            l = l * alpha + ((s0 + s1) + (s2 + s3));
            tmem_stN<PC>(tS + (uint32_t)(hc * PC), pk);
            tmem_st_wait();
          } else {
            l = l * alpha;
          }
          if (g > 0) {
            // Phase bookkeeping: every thread observes every o_full phase in order, and cannot arrive on p_full for
            // tile g before the p_full phase of tile g-1 has completed (PV(g-1) is issued only after it).
            mbar_wait(&o_full, (g - 1) & 1);
          }
          if (tid == 0) FFB_TR(2, 3000 + g);
          tc_fence_before();
          mbar_arrive(&p_full); // this thread's part of P (and of the corrected O) is in place
        }
        // The next segment's Q goes to tensor memory BEFORE this segment's epilogue, so its first QK^T MMAs run
        // while the output of this one is read out (its first PV MMA waits for these threads' p_full anyway).
        if (have) {
          publish_rows();
        }
        // ---- segment epilogue: row sum over the column quarters, then output or split record ----
        int const g_last = g0 + ntiles - 1;
        if (tid == 0) FFB_TR(2, 4000);
        s_xch[0][hc][row] = l;
        mbar_wait(&o_full, g_last & 1);
        if (tid == 0) FFB_TR(2, 4010);
        tc_fence_after();
        asm volatile("bar.sync %0, %1;" ::"r"(quad_bar), "n"(32 * ATC_NQ) : "memory");
        if (tid == 0) FFB_TR(2, 4020);
        l = 0.f;
#pragma unroll
        for (int o = 0; o < ATC_NQ; o++) { // the same order in every thread of the row: one value of l per row
          l += s_xch[0][o][row];
        }
        // registers -> this warp's staging tile (thread per row) -> global with QCH lanes per row, so that a warp
        // store covers 32 / QCH rows x (DH / 2) contiguous bytes (a thread-per-row store touches 32 lines per
        // instruction: LSU-bound).  `f` holds this thread's OC output values.
        auto store_row_slice = [&](float const (&f)[OC]) {
          uint8_t *const mine = wblk + (size_t)lane * (QCH * 16);
#pragma unroll
          for (int c = 0; c < QCH; c++) {
            float v8[8];
#pragma unroll
            for (int e = 0; e < 8; e++) {
              v8[e] = f[c * 8 + e];
            }
            *reinterpret_cast<uint4 *>(mine + ((c ^ swz) << 4)) = pack8<T>(v8);
          }
          __syncwarp();
          constexpr int RPI = 32 / QCH; // rows per store instruction
#pragma unroll
          for (int i = 0; i < QCH; i++) {
            int const rr = i * RPI + lane / QCH, j = lane % QCH;
            int const rswz = QCH == 4 ? ((rr >> 1) & 3) : ((rr >> 2) & 1);
            uint4 const v = *reinterpret_cast<uint4 const *>(wblk + (size_t)rr * (QCH * 16) + ((j ^ rswz) << 4));
            int const thr = __shfl_sync(0xffffffffu, th, rr);
            if (thr >= 0) {
              *reinterpret_cast<uint4 *>(out + (size_t)thr * DH + hc * OC + j * 8) = v;
            }
          }
          __syncwarp();
        };
        if (npieces == 1) {
          if (warp_active) {
            float const inv = __fdividef(1.f, l + 1.e-6f);
            uint32_t r[OC];
            tmem_ldN<OC>(tO, r);
            float f[OC];
#pragma unroll
            for (int e = 0; e < OC; e++) {
              f[e] = __uint_as_float(r[e]) * inv;
            }
            store_row_slice(f);
          }
        } else {
          // Split records live in per-CTA slots laid out [DH + 2][128 rows] (column-major over the rows), so that a
          // warp - one row per lane - writes and later reads 128 contiguous bytes per instruction.  Piece i of a
          // unit is held by CTA c_first + i: slot 2c + 1 when the unit runs past that CTA's range (its last
          // segment), slot 2c when the unit ends there (its first segment).
          constexpr int SLOT = (DH + 2) * TBM;
          if (tid == 0) FFB_TR(2, 4050);
          // Fast path.  Piece 0 of a unit is its CTA's LAST segment, piece 1 (of two) its CTA's FIRST: piece 1's record
          // has nearly always landed by now.  If so (per row), merge straight from tensor memory: no record, no
          // fence, no atomic, one round trip to L2 instead of four.  mode: 0 = general path, 2 = fast, 3 = idle row.
          int mode = 0;
          if (piece == 0 && npieces == 2) {
            if (hc == 0) {
              // the decision is per WARP (32 rows; the tcgen05.ld below needs the warp converged)
              int seen = 1;
              if (th >= 0) {
                asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(split_counter + th) : "memory");
              }
              int const f = __all_sync(0xffffffffu, seen == 1) ? (th >= 0 ? 2 : 3) : 0;
              if (f == 2) {
                split_counter[th] = 0;
              }
              s_rowlast[row] = f;
            }
            asm volatile("bar.sync %0, %1;" ::"r"(quad_bar), "n"(32 * ATC_NQ) : "memory");
            mode = s_rowlast[row];
          }
          if (mode != 0 && warp_active) {
            __threadfence();
            // (idle rows read their - unused - record row too: the loads stay unconditional and batched)
            float const *rs = partial + row + (size_t)(2 * (c_first + 1)) * SLOT;
            float const mb = __ldcg(rs + DH * TBM), lb = __ldcg(rs + (DH + 1) * TBM);
            float vb[OC];
#pragma unroll
            for (int d = 0; d < OC; d++) {
              vb[d] = __ldcg(rs + (hc * OC + d) * TBM);
            }
            // same operations in the same order as the general merge below: bit-identical whichever path runs
            float M = -INFINITY, L = 0.f;
            M = fmaxf(M, m_ref);
            M = fmaxf(M, mb);
            float const wa = m_ref == -INFINITY ? 0.f : ffb_ex2(m_ref - M);
            float const wb = mb == -INFINITY ? 0.f : ffb_ex2(mb - M);
            L += l * wa;
            L += lb * wb;
            float const inv = __fdividef(1.f, L + 1.e-6f);
            uint32_t r[OC];
            tmem_ldN<OC>(tO, r);
            float f[OC];
#pragma unroll
            for (int e = 0; e < OC; e++) {
              f[e] = fmaf(__uint_as_float(r[e]), wa, fmaf(vb[e], wb, 0.f)) * inv;
            }
            if (mode == 2) {
              uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)th * DH + hc * OC);
#pragma unroll
              for (int c = 0; c < QCH; c++) {
                float v8[8];
#pragma unroll
                for (int e = 0; e < 8; e++) {
                  v8[e] = f[c * 8 + e];
                }
                dst[c] = pack8<T>(v8);
              }
            }
          }
          if (mode == 0 && warp_active) {
            float *rec = partial + (size_t)(2 * (int)blockIdx.x + (piece < npieces - 1 ? 1 : 0)) * SLOT + row;
            uint32_t r[OC];
            tmem_ldN<OC>(tO, r);
#pragma unroll
            for (int e = 0; e < OC; e++) {
              rec[(hc * OC + e) * TBM] = __uint_as_float(r[e]);
            }
            if (hc == 0) {
              rec[DH * TBM] = m_ref;
              rec[(DH + 1) * TBM] = l;
            }
          }
          // per-row arrival counters: the last piece to finish a (token, head) row merges the records; every thread
          // of the row merges its own quarter of the columns (exact log-sum-exp merge, then the reference's
          // 1/(sum + 1e-6), inc_…cu:538)
          if (tid == 0) FFB_TR(2, 4100);
          if (mode == 0) {
            __threadfence();
          }
          if (tid == 0) FFB_TR(2, 4200);
          asm volatile("bar.sync %0, %1;" ::"r"(quad_bar), "n"(32 * ATC_NQ) : "memory");
          if (hc == 0 && mode == 0) {
            int last = 0;
            if (th >= 0) {
              int const prev = atomicAdd(&split_counter[th], 1);
              last = prev == npieces - 1;
              if (last) {
                split_counter[th] = 0;
              }
            }
            s_rowlast[row] = last;
          }
          asm volatile("bar.sync %0, %1;" ::"r"(quad_bar), "n"(32 * ATC_NQ) : "memory");
          if (tid == 0) FFB_TR(2, 4300 + s_rowlast[row]);
          if (mode == 0 && s_rowlast[row]) {
            __threadfence();
            float const *rec0 = partial + row;
            auto slot_of = [&](int i) { return rec0 + (size_t)(2 * (c_first + i) + (i < npieces - 1 ? 1 : 0)) * SLOT; };
            float M = -INFINITY, L = 0.f;
            for (int i = 0; i < npieces; i++) {
              M = fmaxf(M, __ldcg(slot_of(i) + DH * TBM));
            }
            for (int i = 0; i < npieces; i++) {
              float const mi = __ldcg(slot_of(i) + DH * TBM);
              L += __ldcg(slot_of(i) + (DH + 1) * TBM) * (mi == -INFINITY ? 0.f : ffb_ex2(mi - M));
            }
            float const inv = __fdividef(1.f, L + 1.e-6f);
            uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)th * DH + hc * OC);
            float acc[OC];
#pragma unroll
            for (int d = 0; d < OC; d++) {
              acc[d] = 0.f;
            }
            // OC columns x 2 pieces per batch: up to 64 independent loads in flight per thread
#pragma unroll 1
            for (int i = 0; i < npieces; i += 2) {
              bool const two = i + 1 < npieces;
              float const *ra = slot_of(i), *rb = slot_of(two ? i + 1 : i);
              float const ma = __ldcg(ra + DH * TBM), mb = __ldcg(rb + DH * TBM);
              float va[OC], vb[OC];
#pragma unroll
              for (int d = 0; d < OC; d++) {
                va[d] = __ldcg(ra + (hc * OC + d) * TBM);
                vb[d] = __ldcg(rb + (hc * OC + d) * TBM);
              }
              float const wa = ma == -INFINITY ? 0.f : ffb_ex2(ma - M);
              float const wb = (!two || mb == -INFINITY) ? 0.f : ffb_ex2(mb - M);
#pragma unroll
              for (int d = 0; d < OC; d++) {
                acc[d] = fmaf(va[d], wa, fmaf(vb[d], wb, acc[d]));
              }
            }
#pragma unroll
            for (int c = 0; c < QCH; c++) {
              float v8[8];
#pragma unroll
              for (int e = 0; e < 8; e++) {
                v8[e] = acc[c * 8 + e] * inv;
              }
              dst[c] = pack8<T>(v8);
            }
            if (tid == 0) FFB_TR(2, 4400);
          }
        }
        if (tid == 0) FFB_TR(2, 5000);
        g0 += ntiles;
      }
      tc_fence_before();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == ATC_MMA_WARP) {
    tc_fence_after();
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base_smem), "r"(512u));
  }
#ifdef FFB_ATC_TRACE_BUILD
  if (trace && tid == 0) {
    long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    trace[1536 + 3 * blockIdx.x + 1] = t;
  }
#endif
}

} // namespace

int attn_tc_supported(ffb_attn_s *h) {
  return h->tm_ready && !ffb::env("FFB_DISABLE_ATTN_TC");
}

int attn_tc_init(ffb_attn_s *h) {
  ffb_attn_desc const &d = h->d;
  long const rows = (long)d.max_requests * d.num_kv_heads * h->cache_len;
  int rc = make_tensor_map_2d(reinterpret_cast<CUtensorMap *>(h->tmK), d.dtype, h->k_cache, rows, d.head_dim, TBN);
  if (rc != FFB_OK) {
    return rc;
  }
  rc = make_tensor_map_2d(reinterpret_cast<CUtensorMap *>(h->tmV), d.dtype, h->v_cache, rows, d.head_dim, TBN);
  if (rc != FFB_OK) {
    return rc;
  }
  h->tm_ready = true;
  return FFB_OK;
}

// Returns FFB_OK when launched, an FFB_ERR_* code on failure, or -1 when the batch does not fit the schedule
// table (the caller then uses the mma.sync kernel).
int launch_attn_tc(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T, AtcReq const *reqs,
                   int nreq, cudaStream_t s) {
  ffb_attn_desc const &d = h->d;
  if (T <= 0 || nreq <= 0) {
    return FFB_OK;
  }
  int const G = d.num_q_heads / d.num_kv_heads;
  // The schedule depends on the batch descriptor and the handle's shape only, and every layer of a step asks for the
  // same one: the search below (~100 candidate range lengths x a walk over all CTAs and units: 40-80 us of host time)
  // runs once per step instead of once per layer.  Per thread, one entry.
  struct SchedKey {
    int nreq, nkv, G, cache_len, mode, head_dim, rg_env;
    size_t partial_bytes;
    AtcReq reqs[FFB_MAX_NUM_REQUESTS];
  };
  struct SchedCache {
    bool valid = false;
    SchedKey key;
    AtcSched sc;
    int cum = 0;
  };
  static thread_local SchedCache cache;
  SchedKey key;
  memset(&key, 0, sizeof(key));
  key.nreq = nreq;
  key.nkv = d.num_kv_heads;
  key.G = G;
  key.cache_len = h->cache_len;
  key.mode = d.mode;
  key.head_dim = d.head_dim;
  key.rg_env = ffb::env("FFB_ATC_RG") ? atoi(ffb::env("FFB_ATC_RG")) : 0;
  key.partial_bytes = h->partial_bytes;
  bool const cacheable = nreq <= FFB_MAX_NUM_REQUESTS;
  if (cacheable) {
    memcpy(key.reqs, reqs, sizeof(AtcReq) * nreq);
  }
  bool const hit = cacheable && cache.valid && memcmp(&cache.key, &key, sizeof(key)) == 0;
  AtcSched sc;
  int cum = 0;
  if (hit) {
    sc = cache.sc;
    cum = cache.cum;
  } else {
  sc.n_ent = 0;
  for (int i = 0; i < nreq; i++) {
    AtcReq const &rq = reqs[i];
    if (rq.qlen <= 0) {
      continue;
    }
    int const m_tiles = (rq.qlen * G + TBM - 1) / TBM;
    for (int mt = 0; mt < m_tiles; mt++) {
      if (sc.n_ent == ATC_MAX_ENT) {
        return -1;
      }
      int const qi_last = std::min(rq.qlen - 1, (mt * TBM + TBM - 1) / G);
      int kv_hi = rq.causal ? std::min(rq.tl, rq.first_depth + qi_last + 1) : rq.tl; // attn_prep.cu kv_len rules
      kv_hi = std::max(1, std::min(kv_hi, h->cache_len));
      int const ntiles = (kv_hi + TBN - 1) / TBN;
      if (ntiles > 32767 || rq.qlen > 32767) {
        return -1;
      }
      AtcEntry &e = sc.ent[sc.n_ent++];
      e.cum = cum;
      e.first_tok = rq.first_tok;
      e.ntcs = rq.ntcs;
      e.slot = (short)rq.slot;
      e.mt = (short)mt;
      e.ntiles = (short)ntiles;
      e.qlen = (short)rq.qlen;
      // beam mode places new rows by the bit-mask sizes (spec_…cu:343-346), not at first_depth: no early streaming there
      e.first_new = d.mode == FFB_BEAM_SEARCH_MODE ? 0 : std::max(0, rq.first_depth);
      cum += d.num_kv_heads * ntiles;
    }
  }
  if (sc.n_ent == 0) {
    return FFB_OK;
  }
  sc.total = cum;
  // Tiles per CTA.  rg0 = ceil(total / SMs) balances perfectly but cuts units at almost every CTA boundary, and a
  // cut costs a split record + merge (~4 tiles of time); a slightly longer range whose boundaries fall on unit
  // boundaries can be cheaper.  Cost model: rg + 4 * (most partial segments any CTA has), rg in [rg0, 2 rg0].
  int const rg0 = (cum + kNumSMs - 1) / kNumSMs;
  auto max_partials = [&](int rgc) {
    int e = 0, worst = 0, prev_cut = 0;
    for (int x = rgc; x < cum + rgc; x += rgc) { // x: end of a CTA range (the last one is clipped to total)
      int cut = 0;
      if (x < cum) {
        while (e + 1 < sc.n_ent && sc.ent[e + 1].cum <= x) {
          e++;
        }
        cut = ((x - sc.ent[e].cum) % sc.ent[e].ntiles) != 0;
      }
      worst = std::max(worst, prev_cut + cut);
      prev_cut = cut;
    }
    return worst;
  };
  // A unit cut into three or more pieces takes the general (record + counter + last-arriver) merge, a dependent chain
  // of L2 round trips: one more penalty per extra piece.  The search runs up to the longest unit, so that a batch
  // with fewer units than SMs (head-sharded shards: 32 units of 9 tiles at TP = 8) runs one whole unit per CTA
  // instead of cutting every unit to fill all SMs (measured 39 us -> see profiles/).
  int max_nt = 1;
  for (int i = 0; i < sc.n_ent; i++) {
    max_nt = std::max(max_nt, (int)sc.ent[i].ntiles);
  }
  auto max_pieces = [&](int rgc) {
    int worst = 1;
    for (int i = 0; i < sc.n_ent; i++) {
      for (int hh = 0; hh < d.num_kv_heads; hh++) {
        int const u0 = sc.ent[i].cum + hh * sc.ent[i].ntiles, u1 = u0 + sc.ent[i].ntiles - 1;
        worst = std::max(worst, u1 / rgc - u0 / rgc + 1);
      }
    }
    return worst;
  };
  int rg = rg0, best_cost = 1 << 30;
  int const rg_hi = std::max(2 * rg0, max_nt), step = std::max(1, (rg_hi - rg0) / 96);
  for (int rgc = rg0; rgc <= rg_hi; rgc += step) {
    int const cost = rgc + 4 * max_partials(rgc) + 4 * std::max(0, max_pieces(rgc) - 2);
    if (cost < best_cost) {
      best_cost = cost;
      rg = rgc;
    }
  }
  if (char const *e = ffb::env("FFB_ATC_RG")) {
    int v = atoi(e);
    if (v > 0) {
      rg = v;
    }
  }
  // split records: two slots of [Dh + 2][128] floats per CTA (see the kernel's epilogue)
  size_t const slot_bytes = (size_t)(d.head_dim + 2) * TBM * sizeof(float);
  while ((size_t)2 * ((cum + rg - 1) / rg) * slot_bytes > h->partial_bytes) {
    rg *= 2; // fewer, longer CTA ranges
  }
  int maxp = 1;
  for (int i = 0; i < sc.n_ent; i++) {
    maxp = std::max(maxp, (sc.ent[i].ntiles - 1 + rg - 1) / rg + 1);
  }
  sc.rg = rg;
  sc.maxpieces = maxp;
  if (cacheable) {
    cache.key = key;
    cache.sc = sc;
    cache.cum = cum;
    cache.valid = true;
  }
  } // !hit
  int const rg = sc.rg, maxp = sc.maxpieces;
  (void)maxp;
  int const grid = (cum + rg - 1) / rg;
  float const scale = d.qk_prod_scaling ? 1.0f / sqrtf((float)d.head_dim) : 1.0f;
  float const scale_log2 = scale * 1.4426950408889634f;
  static long long *trace = nullptr;
  static int trace_calls = 0;
  bool const tracing = ffb::env("FFB_ATC_TRACE") != nullptr;
  if (tracing && !trace) {
    cudaMalloc(&trace, (3 * 512 + 3 * 160) * sizeof(long long));
  }
  if (tracing) {
    cudaMemsetAsync(trace, 0, (3 * 512 + 3 * 160) * sizeof(long long), s);
  }
  CUtensorMap const &tmK = *reinterpret_cast<CUtensorMap const *>(h->tmK);
  CUtensorMap const &tmV = *reinterpret_cast<CUtensorMap const *>(h->tmV);
  if (h->profiling == 1) {
    cudaEventRecord(h->ev0, s);
  }
  // launched as a programmatic dependent of the preceding kernel in the stream (the prep kernel, which signals
  // griddepcontrol.launch_dependents at once): set-up and the stream of old K/V tiles overlap its execution.
  // FFB_ATC_PDL=0 launches it the ordinary way.
  static bool const pdl = !(env("FFB_ATC_PDL") && env("FFB_ATC_PDL")[0] == '0');
  // K / V tiles loaded with the L2 evict-first policy (FFB_ATC_EVICT=0: default policy)
  int const kv_evict = !(env("FFB_ATC_EVICT") && env("FFB_ATC_EVICT")[0] == '0');
  // the K / V stream waits (after its first tile) until the first segment's Q loads have been issued (FFB_ATC_QFIRST=0: no)
  int const q_first = !(env("FFB_ATC_QFIRST") && env("FFB_ATC_QFIRST")[0] == '0');
#define FFB_ATC(TT, DH)                                                                                              \
  do {                                                                                                               \
    size_t const smem = AtcCfg<DH>::SMEM;                                                                            \
    FFB_SET_MAX_SMEM_ONCE((attn_tc_kernel<TT, DH>), smem);                                                           \
    cudaLaunchConfig_t cfg = {};                                                                                     \
    cfg.gridDim = dim3(grid);                                                                                        \
    cfg.blockDim = dim3(ATC_THREADS);                                                                                \
    cfg.dynamicSmemBytes = smem;                                                                                     \
    cfg.stream = s;                                                                                                  \
    cudaLaunchAttribute attr[1];                                                                                     \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                                 \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                                          \
    cfg.attrs = attr;                                                                                                \
    cfg.numAttrs = pdl && h->profiling != 2 ? 1 : 0;                                                                 \
    FFB_CUDA_TRY(cudaLaunchKernelEx(&cfg, attn_tc_kernel<TT, DH>, tmK, tmV, sc, (TT const *)h->q_buf,                \
                                    (QueryDesc const *)h->qdesc, meta, h->partial, h->split_counter, (TT *)out,      \
                                    d.num_q_heads, d.num_kv_heads, h->cache_len, scale_log2, kv_evict, q_first,      \
                                    tracing ? trace : (long long *)nullptr));                                        \
  } while (0)
  if (d.dtype == FFB_DT_HALF) {
    if (d.head_dim == 128) FFB_ATC(__half, 128); else FFB_ATC(__half, 64);
  } else {
    if (d.head_dim == 128) FFB_ATC(__nv_bfloat16, 128); else FFB_ATC(__nv_bfloat16, 64);
  }
#undef FFB_ATC
  if (h->profiling == 1) {
    cudaEventRecord(h->ev1, s);
  }
  FFB_LAUNCH_CHECK();
  if (tracing && ++trace_calls == atoi(ffb::env("FFB_ATC_TRACE"))) { // dump the n-th launch
    static long long host[3 * 512 + 3 * 160];
    cudaStreamSynchronize(s);
    cudaMemcpy(host, trace, sizeof(host), cudaMemcpyDeviceToHost);
    long long t0 = host[1];
    for (int r = 0; r < 3; r++) {
      for (int i = 0; i < 255 && host[r * 512 + 2 * i + 1]; i++) {
        fprintf(stderr, "TRACE role %d ev %lld t %lld\n", r, host[r * 512 + 2 * i], host[r * 512 + 2 * i + 1] - t0);
      }
    }
    for (int c = 0; c < grid; c++) {
      fprintf(stderr, "TRACE cta %d sm %lld start %lld end %lld\n", c, host[1536 + 3 * c + 2], host[1536 + 3 * c] - t0,
              host[1536 + 3 * c + 1] - t0);
    }
    fprintf(stderr, "TRACE grid %d rg %d maxp %d total %d\n", grid, rg, maxp, cum);
  }
  return FFB_OK;
}

} // namespace ffb
