// attn_tc.cu — tcgen05 / TMA multi-query attention over a shared KV stream: tree-verify (bit mask), beam / spec
// (bit mask with the SSM's query-bit rule), chunked-prefill prompt (causal) and GQA decode.
// Replaces compute_attention_kernel_fused_kernel (src/ops/tree_inc_multihead_self_attention.cu:35-333),
// compute_spec_inc_attention_kernel_generation_kernel (src/ops/spec_inc_multihead_self_attention.cu:36-309)
// and the cuBLAS + cuDNN prompt path (src/ops/inc_multihead_self_attention.cu:98-366).
//
// One CTA owns a 128-row query tile (query tokens x the q heads sharing one kv head) of one request and a
// key range of the split; every K/V tile of 128 keys is read from HBM exactly once:
//   warp 8 (one lane)  TMA producer: K and V tiles straight from the [request][kv_head][cache_len][Dh] cache into
//                      128B-swizzled shared-memory rings (separate full/empty mbarriers for K and V)
//   warp 9 (one lane)  MMA issuer: S = Q K^T (UMMA 128x128x16, both operands K-major) into one of two TMEM
//                      score buffers, then O += P V (V consumed MN-major, exactly as TMA laid it down) into the
//                      TMEM output accumulator; tcgen05.commit releases ring slots
//   warps 0-7          two threads per query row (warps w and w+4 share TMEM lane quarter w, each takes 64 of
//                      the 128 keys and half of the output columns): tcgen05.ld the scores, scale + kv_len /
//                      causal / tree-bit mask, online softmax in the log2 domain against a LAZY reference
//                      maximum (the TMEM accumulator is rescaled only when a row maximum grows by more than
//                      2^8, so P stays <= 256 and most tiles need no correction), P written 16-bit into
//                      swizzled smem (A operand of the second MMA)
// The softmax of tile j overlaps the QK^T MMA of tile j+1; the kernel is HBM-bound: 64 KB of K/V
// per tile per SM (~2900 cycles at the HBM rate) against ~1000 tensor-pipe and ~1000 MUFU cycles.
#include "attn_state.h"
#include "tc_common.cuh"

namespace ffb {

int launch_combine(ffb_attn_s *h, void *out, int T, int chunk, int nsplit, cudaStream_t s);

using namespace tc;

namespace {

constexpr int TBM = 128, TBN = 128, ATC_SOFTMAX_THREADS = 256, ATC_THREADS = 320;
constexpr float kRescaleThreshold = 8.f; // log2 domain

template <int DH> struct AtcCfg {
  static constexpr int KH = DH / 64;                 // 64-element (128-byte) column blocks per row
  static constexpr int STAGES = DH == 128 ? 2 : 4;   // K and V ring depth
  static constexpr uint32_t HALF_BYTES = 128 * 128;  // one [128 rows][64 cols] swizzled block
  static constexpr uint32_t Q_BYTES = KH * HALF_BYTES;
  static constexpr uint32_t P_BYTES = 2 * HALF_BYTES;
  static constexpr uint32_t KV_BYTES = KH * HALF_BYTES;
  static constexpr size_t SMEM = Q_BYTES + P_BYTES + 2 * STAGES * KV_BYTES + 1024;
};

template <typename T, int DH>
__global__ void __launch_bounds__(ATC_THREADS, 1)
attn_tc_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
               T const *__restrict__ q_buf, QueryDesc const *__restrict__ qdesc,
               ffb_combined_batch_config_meta const *__restrict__ meta, float *__restrict__ partial,
               int *__restrict__ split_counter, T *__restrict__ out, int mode, int nq, int nkv, int cache_len,
               int m_tiles, int chunk, int nsplit, float scale_log2, int num_generation_tokens, int fuse_combine) {
  using Cfg = AtcCfg<DH>;
  constexpr int KH = Cfg::KH, STAGES = Cfg::STAGES;
  constexpr uint32_t HALF = Cfg::HALF_BYTES;
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t k_full[STAGES], k_empty[STAGES], v_full[STAGES], v_empty[STAGES];
  __shared__ __align__(8) uint64_t s_full[2], s_empty[2], o_full, p_full;
  __shared__ float s_xch[2][2][TBM]; // [tile parity][column half][row]: partial row maxima / final row sums
  __shared__ uint32_t tmem_base_smem;
  __shared__ int s_kvlen[TBM], s_bit[TBM], s_tok[TBM], s_rowlast[TBM];
  __shared__ unsigned long long s_mask[64];
  __shared__ int s_hi, s_lo, s_anybit;

  int const slot = blockIdx.z;
  if (meta->request_completed[slot]) {
    return;
  }
  ffb_per_request_info const *ri = &meta->requestsInfo[slot];
  int const first_tok = ri->first_token_offset_in_batch;
  int qlen = ri->num_tokens_in_batch;
  if (mode == FFB_BEAM_SEARCH_MODE && first_tok < num_generation_tokens) {
    qlen = meta->beamRequestsInfo[slot].sub_request_num; // spec_…cu:97-98
  }
  int const G = nq / nkv;
  int const kvh = blockIdx.y / m_tiles, mt = blockIdx.y % m_tiles;
  int const rows_total = qlen * G;
  if (mt * TBM >= rows_total) {
    return;
  }
  int const split = blockIdx.x, k0 = split * chunk;
  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const ntcs = meta->causalMask[slot].non_tree_cache_size;

  if (tid == 0) {
    s_hi = 0;
    s_lo = 0x7fffffff;
    s_anybit = 0;
  }
  if (tid >= 256) {
    s_mask[tid - 256] = meta->causalMask[slot].mask[tid - 256];
  }
  __syncthreads();
  if (tid < TBM) {
    int const R = mt * TBM + tid;
    int tokhead = -1, kvl = 0, bit = -1;
    if (R < rows_total) {
      int const tok = first_tok + R / G;
      tokhead = tok * nq + kvh * G + R % G;
      QueryDesc const qd = qdesc[tok];
      kvl = qd.kv_len;
      bit = qd.bit;
      atomicMax(&s_hi, kvl);
      atomicMin(&s_lo, kvl);
      if (bit >= 0) {
        atomicOr(&s_anybit, 1);
      }
    }
    s_tok[tid] = tokhead;
    s_kvlen[tid] = kvl;
    s_bit[tid] = bit;
  }
  __syncthreads();
  int const kv_hi = s_hi, kv_lo = s_lo;
  bool const anybit = s_anybit != 0;
  if (split == 0 && kv_lo <= 0 && tid < TBM) { // rows without any key: zero output (no split will touch them)
    if (s_tok[tid] >= 0 && s_kvlen[tid] <= 0) {
      uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)s_tok[tid] * DH);
      for (int c = 0; c < DH / 8; c++) {
        dst[c] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  }
  if (k0 >= kv_hi) {
    return;
  }
  int const k1 = min(k0 + chunk, kv_hi);
  int const ntiles = (k1 - k0 + TBN - 1) / TBN;

  // 128B-swizzled operand blocks need 1024-byte alignment
  uint8_t *smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t *sQ = smem, *sP = sQ + Cfg::Q_BYTES, *sK = sP + Cfg::P_BYTES, *sV = sK + STAGES * Cfg::KV_BYTES;

  if (warp == 9 && lane == 0) {
    for (int s = 0; s < STAGES; s++) {
      mbar_init(&k_full[s], 1);
      mbar_init(&k_empty[s], 1);
      mbar_init(&v_full[s], 1);
      mbar_init(&v_empty[s], 1);
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(&s_full[b], 1);
      mbar_init(&s_empty[b], ATC_SOFTMAX_THREADS);
    }
    mbar_init(&o_full, 1);
    mbar_init(&p_full, ATC_SOFTMAX_THREADS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) { // TMEM: 2 score buffers + the output accumulator, 128 fp32 columns each (512 allocated)
    if (lane == 0) {
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmK) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"(&tmV) : "memory");
    }
    __syncwarp();
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_smem)), "r"(512u));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
  }
  if (tid < TBM) { // Q tile: row `tid`, zero-filled past the request's queries, written in the swizzled K-major layout
    int const th = s_tok[tid];
    uint4 const *src = reinterpret_cast<uint4 const *>(q_buf + (size_t)(th >= 0 ? th : 0) * DH);
#pragma unroll
    for (int c = 0; c < DH / 8; c++) {
      uint4 const v = th >= 0 ? src[c] : make_uint4(0u, 0u, 0u, 0u);
      *reinterpret_cast<uint4 *>(sQ + (c >> 3) * HALF + tid * 128 + (((c & 7) ^ (tid & 7)) << 4)) = v;
    }
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  uint32_t const tmem_base = tmem_base_smem;
  uint32_t const tmem_S = tmem_base, tmem_O = tmem_base + 256;
  int const row_base = (slot * nkv + kvh) * cache_len; // first cache row of this (request, kv head)

  if (warp == 8) {
    if (lane == 0) {
      // ===== TMA producer =====
      for (int it = 0; it < ntiles; it++) {
        int const s = it % STAGES;
        uint32_t const ph = (it / STAGES) & 1;
        int const r0 = row_base + k0 + it * TBN;
        mbar_wait(&k_empty[s], ph ^ 1);
        mbar_expect_tx(&k_full[s], Cfg::KV_BYTES);
#pragma unroll
        for (int hh = 0; hh < KH; hh++) {
          tma_load_2d(sK + s * Cfg::KV_BYTES + hh * HALF, &tmK, hh * 64, r0, &k_full[s]);
        }
        mbar_wait(&v_empty[s], ph ^ 1);
        mbar_expect_tx(&v_full[s], Cfg::KV_BYTES);
#pragma unroll
        for (int hh = 0; hh < KH; hh++) {
          tma_load_2d(sV + s * Cfg::KV_BYTES + hh * HALF, &tmV, hh * 64, r0, &v_full[s]);
        }
      }
    }
  } else if (warp == 9) {
    if (lane == 0) {
      // ===== MMA issuer =====
      // cute::UMMA::InstrDescriptor: c_format F32 (1) @4, a/b format @7/@10 (F16 0, BF16 1), b_major @16 (1 = MN-major),
      // N>>3 @17, M>>4 @24
      constexpr uint32_t fmt = std::is_same<T, __nv_bfloat16>::value ? 1u : 0u;
      constexpr uint32_t idesc_qk = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(TBN >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);
      constexpr uint32_t idesc_pv = (1u << 4) | (fmt << 7) | (fmt << 10) | (1u << 16) | ((uint32_t)(DH >> 3) << 17) | ((uint32_t)(TBM >> 4) << 24);
      auto issue_qk = [&](int it) {
        int const s = it % STAGES;
        mbar_wait(&k_full[s], (it / STAGES) & 1);
        tc_fence_after();
        uint32_t const kb = smem_u32(sK + s * Cfg::KV_BYTES);
#pragma unroll
        for (int k = 0; k < DH / 16; k++) {
          uint64_t const da = umma_desc_sw128(smem_u32(sQ) + (k >> 2) * HALF) + (uint64_t)((k & 3) * 2);
          uint64_t const db = umma_desc_sw128(kb + (k >> 2) * HALF) + (uint64_t)((k & 3) * 2);
          tc_mma_f16(tmem_S + (uint32_t)(it & 1) * 128u, da, db, idesc_qk, k != 0);
        }
        tc_commit(&k_empty[s]);
        tc_commit(&s_full[it & 1]);
      };
      issue_qk(0);
      if (ntiles > 1) {
        issue_qk(1);
      }
      for (int it = 0; it < ntiles; it++) {
        int const s = it % STAGES;
        mbar_wait(&p_full, it & 1);
        mbar_wait(&v_full[s], (it / STAGES) & 1);
        tc_fence_after();
        uint32_t const vb = smem_u32(sV + s * Cfg::KV_BYTES);
#pragma unroll
        for (int k = 0; k < TBN / 16; k++) {
          uint64_t const da = umma_desc_sw128(smem_u32(sP) + (k >> 2) * HALF) + (uint64_t)((k & 3) * 2);
          uint64_t const db = umma_desc_mn_sw128(vb + k * 2048, HALF, 1024);
          tc_mma_f16(tmem_O, da, db, idesc_pv, (it | k) != 0);
        }
        tc_commit(&v_empty[s]);
        tc_commit(&o_full);
        if (it + 2 < ntiles) {
          mbar_wait(&s_empty[it & 1], (it >> 1) & 1);
          issue_qk(it + 2);
        }
      }
    }
  } else {
    // ===== softmax warps: threads (row, h) with h = warp / 4 own keys [64h, 64h+64) of every tile =====
    int const row = tid & (TBM - 1), hc = tid >> 7;
    int const kvl = s_kvlen[row], bit = s_bit[row];
    int const th = s_tok[row];
    bool const warp_active = __any_sync(0xffffffffu, th >= 0); // a warp of padding rows only keeps the barriers moving
    uint32_t const lane_sel = (uint32_t)((warp & 3) * 32) << 16;
    uint32_t const pair_bar = 1 + (warp & 3); // named barrier shared by warps w and w+4
    constexpr int OC = DH / 2;                // output columns per thread
    uint32_t const tO = tmem_O + lane_sel + (uint32_t)(hc * OC);
    float m_ref = -INFINITY, l = 0.f;
    uint8_t *prow = sP + hc * HALF + row * 128;
    int const rx = row & 7;

    for (int it = 0; it < ntiles; it++) {
      int const b = it & 1;
      int const kbase = k0 + it * TBN + hc * 64; // first key of this thread's 64 columns
      uint32_t const tS = tmem_S + (uint32_t)b * 128u + lane_sel + (uint32_t)(hc * 64);
      bool const need_mask = (kbase + 64 > kv_lo) || (kbase + 64 > k1) || (anybit && kbase + 64 > ntcs);
      float sc[64];
      mbar_wait(&s_full[b], (it >> 1) & 1);
      tc_fence_after();
      float mx0 = -INFINITY, mx1 = -INFINITY, mx2 = -INFINITY, mx3 = -INFINITY;
      if (warp_active) {
        uint32_t r0[32], r1[32];
        tmem_ld32_nowait(tS, r0);
        tmem_ld32_nowait(tS + 32u, r1);
        tmem_ld_wait();
#pragma unroll
        for (int e = 0; e < 32; e++) {
          sc[e] = __uint_as_float(r0[e]) * scale_log2;
          sc[32 + e] = __uint_as_float(r1[e]) * scale_log2;
        }
        if (need_mask) {
#pragma unroll
          for (int e = 0; e < 64; e++) {
            int const key = kbase + e;
            bool vis = key < kvl && key < k1;
            if (vis && bit >= 0 && key >= ntcs) {
              int const idx = key - ntcs;
              vis = idx < 64 && bit < 64 && ((s_mask[idx & 63] >> (bit & 63)) & 1ULL);
            }
            sc[e] = vis ? sc[e] : -INFINITY;
          }
        }
#pragma unroll
        for (int e = 0; e < 64; e += 4) {
          mx0 = fmaxf(mx0, sc[e]);
          mx1 = fmaxf(mx1, sc[e + 1]);
          mx2 = fmaxf(mx2, sc[e + 2]);
          mx3 = fmaxf(mx3, sc[e + 3]);
        }
      }
      float mx = fmaxf(fmaxf(mx0, mx1), fmaxf(mx2, mx3));
      s_xch[b][hc][row] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
      mx = fmaxf(mx, s_xch[b][hc ^ 1][row]);
      // lazy reference maximum: move it (and rescale the accumulator) only when the row maximum outgrows it by 2^8
      bool const grow = mx > m_ref + kRescaleThreshold || (m_ref == -INFINITY && mx > -INFINITY);
      float const alpha = grow ? (m_ref == -INFINITY ? 1.f : ffb_ex2(m_ref - mx)) : 1.f;
      if (grow) {
        m_ref = mx;
      }
      float const base = m_ref == -INFINITY ? 0.f : m_ref;
      if (it > 0) { // the PV MMA of the previous tile has finished: P may be overwritten, O may be corrected
        mbar_wait(&o_full, (it - 1) & 1);
        tc_fence_after();
        if (__any_sync(0xffffffffu, alpha != 1.f)) {
#pragma unroll
          for (int c = 0; c < OC / 32; c++) {
            uint32_t r[32];
            tmem_ld32(tO + (uint32_t)(c * 32), r);
#pragma unroll
            for (int e = 0; e < 32; e++) {
              r[e] = __float_as_uint(__uint_as_float(r[e]) * alpha);
            }
            tmem_st32(tO + (uint32_t)(c * 32), r);
          }
          tmem_st_wait();
        }
      }
      if (warp_active) {
        float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int v4 = 0; v4 < 8; v4++) {
          float f[8];
#pragma unroll
          for (int e = 0; e < 8; e++) {
            f[e] = ffb_ex2(sc[v4 * 8 + e] - base);
          }
          s0 += f[0] + f[4];
          s1 += f[1] + f[5];
          s2 += f[2] + f[6];
          s3 += f[3] + f[7];
          *reinterpret_cast<uint4 *>(prow + ((v4 ^ rx) << 4)) = pack8<T>(f);
        }
        l = l * alpha + ((s0 + s1) + (s2 + s3));
      }
      tc_fence_before();
      mbar_arrive(&s_empty[b]); // score buffer b may be overwritten by QK^T of tile it+2
      fence_proxy_async();
      mbar_arrive(&p_full);     // this thread's part of P (and of the corrected O) is in place
    }
    // row sum over both column halves
    s_xch[0][hc][row] = l;
    mbar_wait(&o_full, (ntiles - 1) & 1);
    tc_fence_after();
    asm volatile("bar.sync %0, 64;" ::"r"(pair_bar) : "memory");
    l += s_xch[0][hc ^ 1][row];
    if (nsplit == 1) {
      if (warp_active) {
        float const inv = __fdividef(1.f, l + 1.e-6f);
        uint4 *dst = reinterpret_cast<uint4 *>(out + (size_t)(th >= 0 ? th : 0) * DH + hc * OC);
#pragma unroll
        for (int c = 0; c < OC / 32; c++) {
          uint32_t r[32];
          tmem_ld32(tO + (uint32_t)(c * 32), r);
          if (th >= 0) {
#pragma unroll
            for (int v4 = 0; v4 < 4; v4++) {
              float f[8];
#pragma unroll
              for (int e = 0; e < 8; e++) {
                f[e] = __uint_as_float(r[v4 * 8 + e]) * inv;
              }
              dst[c * 4 + v4] = pack8<T>(f);
            }
          }
        }
      }
    } else {
      if (warp_active) {
        float *rec = partial + ((size_t)(th >= 0 ? th : 0) * nsplit + split) * (DH + 2);
#pragma unroll
        for (int c = 0; c < OC / 32; c++) {
          uint32_t r[32];
          tmem_ld32(tO + (uint32_t)(c * 32), r);
          if (th >= 0) {
#pragma unroll
            for (int e = 0; e < 32; e += 2) {
              *reinterpret_cast<float2 *>(rec + hc * OC + c * 32 + e) = make_float2(__uint_as_float(r[e]), __uint_as_float(r[e + 1]));
            }
          }
        }
        if (th >= 0 && hc == 0) {
          rec[DH] = m_ref;
          rec[DH + 1] = l;
        }
      }
      if (fuse_combine) {
        // per-row arrival counters: the last split to finish a (token, head) row merges its records
        __threadfence();
        asm volatile("bar.sync 5, 256;" ::: "memory");
        if (hc == 0) {
          int last = 0;
          if (th >= 0 && k0 < kvl) {
            int const nvalid = min(nsplit, (kvl + chunk - 1) / chunk);
            int const prev = atomicAdd(&split_counter[th], 1);
            last = prev == nvalid - 1;
            if (last) {
              split_counter[th] = 0;
            }
          }
          s_rowlast[row] = last;
        }
        asm volatile("bar.sync 5, 256;" ::: "memory");
        __threadfence();
        for (int r = warp; r < TBM; r += 8) {
          if (s_rowlast[r]) {
            int const t2 = s_tok[r];
            int const nvalid = min(nsplit, (s_kvlen[r] + chunk - 1) / chunk);
            combine_record<T, DH>(partial + (size_t)t2 * nsplit * (DH + 2), nvalid, out + (size_t)t2 * DH, lane, 32);
          }
        }
      }
    }
    tc_fence_before();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    tc_fence_after();
    __syncwarp();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512u));
  }
}

} // namespace

// Split choice: CTAs are (request, kv head, row tile, split); one CTA per SM is resident, so pick the chunk
// (a multiple of the 128-key tile) that minimises waves x tiles-per-CTA plus a per-CTA fixed cost.
static int choose_tc_chunk(long units, int kv_len_max, int *nsplit_out) {
  if (char const *e = getenv("FFB_ATTN_TC_CHUNK")) {
    int v = atoi(e);
    if (v > 0) {
      v = ((v + TBN - 1) / TBN) * TBN;
      *nsplit_out = (kv_len_max + v - 1) / v;
      return v;
    }
  }
  int const tiles_total = (kv_len_max + TBN - 1) / TBN;
  double best = 1e30;
  int best_ns = 1, best_chunk = tiles_total * TBN;
  for (int ns = 1; ns <= 16 && ns <= tiles_total; ns++) {
    int const tpc = (tiles_total + ns - 1) / ns; // tiles per chunk
    int const real_ns = (tiles_total + tpc - 1) / tpc;
    if (real_ns != ns) {
      continue;
    }
    long const ctas = units * ns;
    long const waves = (ctas + kNumSMs - 1) / kNumSMs;
    double const cost = (double)waves * (tpc + 1.5) + (ns > 1 ? 0.5 : 0.0);
    if (cost < best - 1e-9) {
      best = cost;
      best_ns = ns;
      best_chunk = tpc * TBN;
    }
  }
  *nsplit_out = best_ns;
  return best_chunk;
}

int attn_tc_supported(ffb_attn_s *h) {
  return h->tm_ready && !getenv("FFB_DISABLE_ATTN_TC");
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

int launch_attn_tc(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T, int kv_len_max,
                   int q_len_max, cudaStream_t s) {
  ffb_attn_desc const &d = h->d;
  if (T <= 0) {
    return FFB_OK;
  }
  int const G = d.num_q_heads / d.num_kv_heads;
  int const m_tiles = (q_len_max * G + TBM - 1) / TBM;
  long const units = (long)((T + q_len_max - 1) / (q_len_max > 0 ? q_len_max : 1)) * d.num_kv_heads * m_tiles;
  int nsplit = 1;
  int chunk = choose_tc_chunk(units, kv_len_max < 1 ? 1 : kv_len_max, &nsplit);
  size_t const rec = (size_t)(d.head_dim + 2) * sizeof(float);
  while (nsplit > 1 && (size_t)T * d.num_q_heads * nsplit * rec > h->partial_bytes) {
    chunk *= 2;
    nsplit = (kv_len_max + chunk - 1) / chunk;
  }
  float const scale = d.qk_prod_scaling ? 1.0f / sqrtf((float)d.head_dim) : 1.0f;
  float const scale_log2 = scale * 1.4426950408889634f;
  int const fuse = getenv("FFB_SEPARATE_COMBINE") ? 0 : 1;
  dim3 grid(nsplit, d.num_kv_heads * m_tiles, d.max_requests);
  CUtensorMap const &tmK = *reinterpret_cast<CUtensorMap const *>(h->tmK);
  CUtensorMap const &tmV = *reinterpret_cast<CUtensorMap const *>(h->tmV);
  if (h->profiling) {
    cudaEventRecord(h->ev0, s);
  }
#define FFB_ATC(TT, DH)                                                                                              \
  do {                                                                                                               \
    size_t const smem = AtcCfg<DH>::SMEM;                                                                            \
    static bool attr_set = false;                                                                                    \
    if (!attr_set) {                                                                                                 \
      FFB_CUDA_TRY(cudaFuncSetAttribute(attn_tc_kernel<TT, DH>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); \
      attr_set = true;                                                                                               \
    }                                                                                                                \
    attn_tc_kernel<TT, DH><<<grid, ATC_THREADS, smem, s>>>(                                                          \
        tmK, tmV, (TT const *)h->q_buf, h->qdesc, meta, h->partial, h->split_counter, (TT *)out, d.mode,             \
        d.num_q_heads, d.num_kv_heads, h->cache_len, m_tiles, chunk, nsplit, scale_log2,                             \
        h->num_generation_tokens_cur, fuse);                                                                         \
  } while (0)
  if (d.dtype == FFB_DT_HALF) {
    if (d.head_dim == 128) FFB_ATC(__half, 128); else FFB_ATC(__half, 64);
  } else {
    if (d.head_dim == 128) FFB_ATC(__nv_bfloat16, 128); else FFB_ATC(__nv_bfloat16, 64);
  }
#undef FFB_ATC
  if (h->profiling) {
    cudaEventRecord(h->ev1, s);
  }
  FFB_LAUNCH_CHECK();
  if (nsplit > 1 && !fuse) {
    return launch_combine(h, out, T, chunk, nsplit, s);
  }
  return FFB_OK;
}

} // namespace ffb
