// attn_multi.cu — multi-query attention over a shared KV stream: tree-verify (bit-mask), beam/spec
// (bit-mask with the SSM's query-bit rule), chunked-prefill prompt (causal) and GQA decode.
// Replaces compute_attention_kernel_fused_kernel (src/ops/tree_inc_multihead_self_attention.cu:35-333),
// compute_spec_inc_attention_kernel_generation_kernel (src/ops/spec_inc_multihead_self_attention.cu:36-309)
// and the cuBLAS+cuDNN prompt path (src/ops/inc_multihead_self_attention.cu:98-366).
// The reference re-streams the whole KV cache once per query token; here all query rows
// of a request (x the q heads sharing a kv head) sit in one 64-row tile and every K/V tile is read
// from HBM once, staged in shared memory with cp.async (double-buffered, XOR-swizzled), multiplied
// on the tensor cores (mma.sync m16n8k16, fp32 accumulate) with an online softmax in registers.
// The kernel is HBM-bound by design; tcgen05 is reserved for the projection GEMMs (linear_*.cu).
#include "attn_state.h"

namespace ffb {

int launch_combine(ffb_attn_s *h, void *out, int T, int chunk, int nsplit, cudaStream_t s);
int choose_chunk(ffb_attn_s *h, long units, int kv_len_max, int granule, long target_ctas,
                 char const *env_name, int *nsplit_out);

namespace {

constexpr int BM = 64, BN = 64, MQ_THREADS = 256; // 8 warps = 4 row groups (16 query rows) x 2 key halves (32 keys)

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t smem_u32(void const *p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, void const *src, int src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;"); }
template <int N> __device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N));
}
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t &r0, uint32_t &r1, uint32_t &r2, uint32_t &r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
template <typename T>
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__half>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                 uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2,
                                                        uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// byte offset of 16-byte chunk c of row r in a [rows][DH] 16-bit tile, XOR-swizzled so the 8 rows
// an ldmatrix touches fall in 8 distinct bank groups
template <int DH> __device__ __forceinline__ uint32_t swz(int r, int c) {
  return (uint32_t)(r * (DH * 2) + ((c ^ (r & 7)) << 4));
}

} // namespace

namespace {

template <typename T, int DH>
__global__ void __launch_bounds__(MQ_THREADS, 2)
multi_attn_kernel(T const *__restrict__ q_buf, T const *__restrict__ k_cache, T const *__restrict__ v_cache,
                  QueryDesc const *__restrict__ qdesc, ffb_combined_batch_config_meta const *__restrict__ meta,
                  float *__restrict__ partial, int *__restrict__ split_counter, T *__restrict__ out, int mode, int nq,
                  int nkv, int cache_len, int m_tiles, int chunk, int nsplit, float scale_log2,
                  int num_generation_tokens, int fuse_combine) {
  constexpr int CPR = DH / 8;            // 16-byte chunks per row
  constexpr int TILE_BYTES = BN * DH * 2;
  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t *sQ = smem;
  uint8_t *sK = smem + BM * DH * 2;
  uint8_t *sV = sK + 2 * TILE_BYTES;
  __shared__ int s_kvlen[BM], s_bit[BM], s_tok[BM];
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
  if (mt * BM >= rows_total) {
    return;
  }
  int const split = blockIdx.x, k0 = split * chunk;
  int const tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int const rg = warp & 3, kh = warp >> 2; // row group, key half of every 64-key tile
  int const ntcs = meta->causalMask[slot].non_tree_cache_size;

  if (tid == 0) {
    s_hi = 0;
    s_lo = 0x7fffffff;
    s_anybit = 0;
  }
  if (tid < 64) {
    s_mask[tid] = meta->causalMask[slot].mask[tid];
  }
  __syncthreads();
  if (tid < BM) {
    int const R = mt * BM + tid;
    bool const valid = R < rows_total;
    int tokhead = -1, kvl = 0, bit = -1;
    if (valid) {
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
  int const rA = rg * 16 + (lane >> 2), rB = rA + 8;
  int const tq = lane & 3;
  if (split == 0 && kv_lo <= 0) { // rows without any key: zero output (no split will touch them)
    for (int i = tid; i < BM * (DH / 2); i += MQ_THREADS) {
      int r = i / (DH / 2), c = i % (DH / 2);
      if (s_tok[r] >= 0 && s_kvlen[r] <= 0) {
        reinterpret_cast<uint32_t *>(out + (size_t)s_tok[r] * DH)[c] = 0u;
      }
    }
  }
  if (k0 >= kv_hi) {
    return;
  }
  int const k1 = min(k0 + chunk, kv_hi);
  int const ntiles = (k1 - k0 + BN - 1) / BN;
  size_t const head_off = ((size_t)slot * nkv + kvh) * cache_len * DH;
  T const *kbase = k_cache + head_off;
  T const *vbase = v_cache + head_off;
  uint32_t const sQa = smem_u32(sQ), sKa = smem_u32(sK), sVa = smem_u32(sV);

  // Q tile (zero-filled for rows past the request's queries)
  for (int i = tid; i < BM * CPR; i += MQ_THREADS) {
    int const r = i / CPR, c = i % CPR;
    int const th = s_tok[r];
    T const *src = th >= 0 ? q_buf + (size_t)th * DH + c * 8 : q_buf;
    cp_async16(sQa + swz<DH>(r, c), src, th >= 0 ? 16 : 0);
  }
  auto load_tile = [&](int it, int buf) {
    int const kt = k0 + it * BN;
    for (int i = tid; i < BN * CPR; i += MQ_THREADS) {
      int const r = i / CPR, c = i % CPR;
      int const key = kt + r;
      bool const ok = key < k1;
      size_t const off = ok ? (size_t)key * DH + c * 8 : 0;
      uint32_t const d = buf * TILE_BYTES + swz<DH>(r, c);
      cp_async16(sKa + d, kbase + off, ok ? 16 : 0);
      cp_async16(sVa + d, vbase + off, ok ? 16 : 0);
    }
  };
  load_tile(0, 0);
  cp_async_commit();

  int const kvlA = s_kvlen[rA], kvlB = s_kvlen[rB];
  int const bitA = s_bit[rA], bitB = s_bit[rB];
  float o[DH / 8][4];
#pragma unroll
  for (int j = 0; j < DH / 8; j++) {
    o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  }
  float mA = -INFINITY, mB = -INFINITY, lA = 0.f, lB = 0.f;

  for (int it = 0; it < ntiles; it++) {
    int const buf = it & 1;
    if (it + 1 < ntiles) {
      load_tile(it + 1, buf ^ 1);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    int const kt = k0 + it * BN;
    uint32_t const sKb = sKa + buf * TILE_BYTES, sVb = sVa + buf * TILE_BYTES;

    // S = Q K^T  (16 rows x 32 keys per warp)
    float s[4][4];
#pragma unroll
    for (int j = 0; j < 4; j++) {
      s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
    }
#pragma unroll
    for (int ks = 0; ks < DH / 16; ks++) {
      uint32_t a0, a1, a2, a3;
      ldsm_x4(sQa + swz<DH>(rg * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, ks * 2 + (lane >> 4)), a0, a1, a2, a3);
#pragma unroll
      for (int np = 0; np < 2; np++) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4(sKb + swz<DH>(kh * 32 + np * 16 + (lane & 7) + (lane >> 4) * 8, ks * 2 + ((lane >> 3) & 1)), b0, b1, b2, b3);
        mma16816<T>(s[2 * np], a0, a1, a2, a3, b0, b1);
        mma16816<T>(s[2 * np + 1], a0, a1, a2, a3, b2, b3);
      }
    }
    // scale, mask (kv_len / causal bound, tree bit mask), online softmax
    bool const need_mask = (kt + BN > kv_lo) || (kt + BN > k1) || (anybit && kt + BN > ntcs);
    float mxA = -INFINITY, mxB = -INFINITY;
#pragma unroll
    for (int j = 0; j < 4; j++) {
#pragma unroll
      for (int e = 0; e < 4; e++) {
        float v = s[j][e] * scale_log2;
        if (need_mask) {
          int const key = kt + kh * 32 + j * 8 + 2 * tq + (e & 1);
          int const kvl = e < 2 ? kvlA : kvlB;
          int const bit = e < 2 ? bitA : bitB;
          bool vis = key < kvl && key < k1;
          if (vis && bit >= 0 && key >= ntcs) {
            int const idx = key - ntcs;
            vis = idx < 64 && bit < 64 && ((s_mask[idx & 63] >> (bit & 63)) & 1ULL);
          }
          v = vis ? v : -INFINITY;
        }
        s[j][e] = v;
        if (e < 2) {
          mxA = fmaxf(mxA, v);
        } else {
          mxB = fmaxf(mxB, v);
        }
      }
    }
    mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 1));
    mxA = fmaxf(mxA, __shfl_xor_sync(0xffffffffu, mxA, 2));
    mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 1));
    mxB = fmaxf(mxB, __shfl_xor_sync(0xffffffffu, mxB, 2));
    float const mAn = fmaxf(mA, mxA), mBn = fmaxf(mB, mxB);
    float const cA = mAn == -INFINITY ? 1.f : ex2(mA - mAn);
    float const cB = mBn == -INFINITY ? 1.f : ex2(mB - mBn);
    float const bA = mAn == -INFINITY ? 0.f : mAn, bB = mBn == -INFINITY ? 0.f : mBn;
    mA = mAn;
    mB = mBn;
    float sumA = 0.f, sumB = 0.f;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      s[j][0] = ex2(s[j][0] - bA);
      s[j][1] = ex2(s[j][1] - bA);
      s[j][2] = ex2(s[j][2] - bB);
      s[j][3] = ex2(s[j][3] - bB);
      sumA += s[j][0] + s[j][1];
      sumB += s[j][2] + s[j][3];
    }
    lA = lA * cA + sumA;
    lB = lB * cB + sumB;
#pragma unroll
    for (int j = 0; j < DH / 8; j++) {
      o[j][0] *= cA;
      o[j][1] *= cA;
      o[j][2] *= cB;
      o[j][3] *= cB;
    }
    // O += P V  (this warp's 32 keys)
#pragma unroll
    for (int kk = 0; kk < 2; kk++) {
      uint32_t const a0 = DT<T>::pack(s[2 * kk][0], s[2 * kk][1]);
      uint32_t const a1 = DT<T>::pack(s[2 * kk][2], s[2 * kk][3]);
      uint32_t const a2 = DT<T>::pack(s[2 * kk + 1][0], s[2 * kk + 1][1]);
      uint32_t const a3 = DT<T>::pack(s[2 * kk + 1][2], s[2 * kk + 1][3]);
#pragma unroll
      for (int dp = 0; dp < DH / 16; dp++) {
        uint32_t b0, b1, b2, b3;
        ldsm_x4_t(sVb + swz<DH>(kh * 32 + kk * 16 + (lane & 7) + ((lane >> 3) & 1) * 8, dp * 2 + (lane >> 4)), b0, b1, b2, b3);
        mma16816<T>(o[2 * dp], a0, a1, a2, a3, b0, b1);
        mma16816<T>(o[2 * dp + 1], a0, a1, a2, a3, b2, b3);
      }
    }
    __syncthreads(); // everyone is done with `buf` before it is refilled
  }

  // merge the two key halves: thread (warp w+4, lane) hands its fragment-identical state to (warp w, lane)
  {
    constexpr int XW = DH / 2 + 4; // floats per thread: o fragment + mA, mB, lA, lB
    float *xch = reinterpret_cast<float *>(sK) + ((size_t)rg * 32 + lane) * XW; // K tiles are dead after the loop's last barrier
    if (kh == 1) {
#pragma unroll
      for (int j = 0; j < DH / 8; j++) {
        *reinterpret_cast<float4 *>(xch + j * 4) = make_float4(o[j][0], o[j][1], o[j][2], o[j][3]);
      }
      *reinterpret_cast<float4 *>(xch + DH / 2) = make_float4(mA, mB, lA, lB);
    }
    __syncthreads();
    if (kh == 1) {
      return; // nothing after this point synchronises the whole block
    }
    float4 const st = *reinterpret_cast<float4 const *>(xch + DH / 2);
    float const mAn = fmaxf(mA, st.x), mBn = fmaxf(mB, st.y);
    float const c1A = mAn == -INFINITY ? 1.f : ex2(mA - mAn), c2A = mAn == -INFINITY ? 1.f : ex2(st.x - mAn);
    float const c1B = mBn == -INFINITY ? 1.f : ex2(mB - mBn), c2B = mBn == -INFINITY ? 1.f : ex2(st.y - mBn);
    lA = lA * c1A + st.z * c2A;
    lB = lB * c1B + st.w * c2B;
    mA = mAn;
    mB = mBn;
#pragma unroll
    for (int j = 0; j < DH / 8; j++) {
      float4 const v = *reinterpret_cast<float4 const *>(xch + j * 4);
      o[j][0] = o[j][0] * c1A + v.x * c2A;
      o[j][1] = o[j][1] * c1A + v.y * c2A;
      o[j][2] = o[j][2] * c1B + v.z * c2B;
      o[j][3] = o[j][3] * c1B + v.w * c2B;
    }
  }
  lA += __shfl_xor_sync(0xffffffffu, lA, 1);
  lA += __shfl_xor_sync(0xffffffffu, lA, 2);
  lB += __shfl_xor_sync(0xffffffffu, lB, 1);
  lB += __shfl_xor_sync(0xffffffffu, lB, 2);
  int const thA = s_tok[rA], thB = s_tok[rB];
  if (nsplit == 1) {
    float const iA = __fdividef(1.f, lA + 1.e-6f), iB = __fdividef(1.f, lB + 1.e-6f);
#pragma unroll
    for (int j = 0; j < DH / 8; j++) {
      if (thA >= 0) {
        reinterpret_cast<uint32_t *>(out + (size_t)thA * DH)[j * 4 + tq] = DT<T>::pack(o[j][0] * iA, o[j][1] * iA);
      }
      if (thB >= 0) {
        reinterpret_cast<uint32_t *>(out + (size_t)thB * DH)[j * 4 + tq] = DT<T>::pack(o[j][2] * iB, o[j][3] * iB);
      }
    }
  } else {
    if (thA >= 0) {
      float *rec = partial + ((size_t)thA * nsplit + split) * (DH + 2);
#pragma unroll
      for (int j = 0; j < DH / 8; j++) {
        *reinterpret_cast<float2 *>(rec + j * 8 + 2 * tq) = make_float2(o[j][0], o[j][1]);
      }
      if (tq == 0) {
        rec[DH] = mA;
        rec[DH + 1] = lA;
      }
    }
    if (thB >= 0) {
      float *rec = partial + ((size_t)thB * nsplit + split) * (DH + 2);
#pragma unroll
      for (int j = 0; j < DH / 8; j++) {
        *reinterpret_cast<float2 *>(rec + j * 8 + 2 * tq) = make_float2(o[j][2], o[j][3]);
      }
      if (tq == 0) {
        rec[DH] = mB;
        rec[DH + 1] = lB;
      }
    }
    if (fuse_combine) {
      // per-row arrival counters: the last split to finish a (token, head) row merges its records
      __shared__ int s_rowlast[BM];
      __threadfence();
      asm volatile("bar.sync 2, 128;" ::: "memory"); // only the 4 surviving warps
      if (tid < BM) {
        int const th = s_tok[tid];
        int last = 0;
        if (th >= 0 && k0 < s_kvlen[tid]) {
          int const nvalid = min(nsplit, (s_kvlen[tid] + chunk - 1) / chunk);
          int const prev = atomicAdd(&split_counter[th], 1);
          last = prev == nvalid - 1;
          if (last) {
            split_counter[th] = 0;
          }
        }
        s_rowlast[tid] = last;
      }
      asm volatile("bar.sync 2, 128;" ::: "memory");
      __threadfence();
      for (int r = warp; r < BM; r += 4) {
        if (s_rowlast[r]) {
          int const th = s_tok[r];
          int const nvalid = min(nsplit, (s_kvlen[r] + chunk - 1) / chunk);
          combine_record<T, DH>(partial + (size_t)th * nsplit * (DH + 2), nvalid, out + (size_t)th * DH, lane, 32);
        }
      }
    }
  }
}

} // namespace

int launch_multi_attn(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T,
                      int kv_len_max, int q_len_max, cudaStream_t s) {
  ffb_attn_desc const &d = h->d;
  if (T <= 0) {
    return FFB_OK;
  }
  int const G = d.num_q_heads / d.num_kv_heads;
  int const m_tiles = (q_len_max * G + BM - 1) / BM;
  int nsplit = 1;
  // units: upper bound on (request, kv head, m tile) blocks; >= 2 waves at 2 blocks/SM
  long const units = (long)((T + q_len_max - 1) / (q_len_max > 0 ? q_len_max : 1)) * d.num_kv_heads * m_tiles;
  int chunk = choose_chunk(h, units, kv_len_max < 1 ? 1 : kv_len_max, BN, (long)kNumSMs * 4,
                           "FFB_MULTI_CHUNK", &nsplit);
  size_t const rec = (size_t)(d.head_dim + 2) * sizeof(float);
  while (nsplit > 1 && (size_t)T * d.num_q_heads * nsplit * rec > h->partial_bytes) {
    chunk *= 2;
    nsplit = (kv_len_max + chunk - 1) / chunk;
  }
  float const scale = d.qk_prod_scaling ? 1.0f / sqrtf((float)d.head_dim) : 1.0f;
  float const scale_log2 = scale * 1.4426950408889634f;
  int const fuse = ffb::env("FFB_SEPARATE_COMBINE") ? 0 : 1;
  dim3 grid(nsplit, d.num_kv_heads * m_tiles, d.max_requests);
  if (h->profiling == 1) {
    cudaEventRecord(h->ev0, s);
  }
#define FFB_MQ(TT, DH)                                                                                     \
  do {                                                                                                     \
    size_t const smem = (size_t)(BM + 4 * BN) * DH * 2;                                                    \
    FFB_SET_MAX_SMEM_ONCE((multi_attn_kernel<TT, DH>), smem);                                              \
    multi_attn_kernel<TT, DH><<<grid, MQ_THREADS, smem, s>>>(                                              \
        (TT const *)h->q_buf, (TT const *)h->k_cache, (TT const *)h->v_cache, h->qdesc, meta, h->partial,  \
        h->split_counter, (TT *)out, d.mode, d.num_q_heads, d.num_kv_heads, h->cache_len, m_tiles, chunk, nsplit, \
        scale_log2, h->num_generation_tokens_cur, fuse);                                                                     \
  } while (0)
  if (d.dtype == FFB_DT_HALF) {
    if (d.head_dim == 128) FFB_MQ(__half, 128); else FFB_MQ(__half, 64);
  } else {
    if (d.head_dim == 128) FFB_MQ(__nv_bfloat16, 128); else FFB_MQ(__nv_bfloat16, 64);
  }
#undef FFB_MQ
  if (h->profiling == 1) {
    cudaEventRecord(h->ev1, s);
  }
  FFB_LAUNCH_CHECK();
  if (nsplit > 1 && !fuse) {
    return launch_combine(h, out, T, chunk, nsplit, s);
  }
  return FFB_OK;
}

} // namespace ffb
