// attn_decode.cu — bandwidth-bound single-query ("generation") attention.
// Replaces compute_attention_kernel_generation_kernel (src/ops/inc_multihead_self_attention.cu:372-623):
// that kernel runs one 128-thread block per (head, request) — 512 blocks for 7B at batch 16 — with
// 8-byte K loads, an L-sized shared logits buffer and serial K-then-V phases.  Here the KV range is
// split across blocks (flash-decoding) so 148 SMs stay full, K and V rows are fetched together as
// 128-bit streaming loads with 2*U loads in flight per thread, softmax is online (no L-sized
// buffer, so no 48 KiB context cap), and a small combine pass applies the reference's
// 1/(sum + 1e-6) normalisation.
#include "attn_state.h"

namespace ffb {

__device__ __forceinline__ float fast_exp2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

constexpr int kDecodeThreads = 128;
constexpr int kDecodeWarps = kDecodeThreads / 32;

// partial record per (token, head, split): DH floats of unnormalised output, then m (log2 domain), l
template <typename T, int DH, int U>
__global__ void __launch_bounds__(kDecodeThreads)
decode_attn_kernel(T const *__restrict__ q_buf, T const *__restrict__ k_cache,
                   T const *__restrict__ v_cache, QueryDesc const *__restrict__ qdesc,
                   float *__restrict__ partial, int *__restrict__ split_counter, T *__restrict__ out, int nq,
                   int nkv, int cache_len, int chunk, int nsplit, float scale_log2, int fuse_combine) {
  constexpr int LPK = DH / 8;   // lanes per key row (16-byte slices)
  constexpr int KPW = 32 / LPK; // keys per warp-wide load
  constexpr int GROUPS = kDecodeWarps * KPW;
  int const split = blockIdx.x, head = blockIdx.y, t = blockIdx.z;
  QueryDesc const qd = qdesc[t];
  int const kv_len = qd.kv_len;
  int const k0 = split * chunk;
  if (k0 >= kv_len) {
    if (kv_len <= 0 && split == 0 && threadIdx.x < DH) {
      out[((size_t)t * nq + head) * DH + threadIdx.x] = DT<T>::from_f(0.f);
    }
    return;
  }
  int const k1 = min(k0 + chunk, kv_len);
  int const kvh = head / (nq / nkv);
  size_t const head_off = ((size_t)qd.slot * nkv + kvh) * cache_len * DH;
  T const *kbase = k_cache + head_off;
  T const *vbase = v_cache + head_off;
  int const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int const sub = lane / LPK, sl = lane % LPK;

  float q[8];
  {
    uint4 uq = *reinterpret_cast<uint4 const *>(q_buf + ((size_t)t * nq + head) * DH + sl * 8);
    unpack8<T>(uq, q);
#pragma unroll
    for (int e = 0; e < 8; e++) {
      q[e] *= scale_log2;
    }
  }
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; e++) {
    acc[e] = 0.f;
  }
  float m = -INFINITY, l = 0.f;

  for (int kb = k0 + warp * (U * KPW); kb < k1; kb += kDecodeWarps * U * KPW) {
    uint4 kr[U], vr[U];
#pragma unroll
    for (int u = 0; u < U; u++) {
      int const key = min(kb + u * KPW + sub, k1 - 1);
      kr[u] = ldg_stream(kbase + (size_t)key * DH + sl * 8);
      vr[u] = ldg_stream(vbase + (size_t)key * DH + sl * 8);
    }
    float s[U];
    float mx = m;
#pragma unroll
    for (int u = 0; u < U; u++) {
      float kf[8];
      unpack8<T>(kr[u], kf);
      float d = 0.f;
#pragma unroll
      for (int e = 0; e < 8; e++) {
        d = fmaf(q[e], kf[e], d);
      }
#pragma unroll
      for (int o = LPK / 2; o >= 1; o >>= 1) {
        d += __shfl_xor_sync(0xffffffffu, d, o);
      }
      s[u] = (kb + u * KPW + sub < k1) ? d : -INFINITY;
      mx = fmaxf(mx, s[u]);
    }
    // branch-free update (a branch here lets the compiler sink the V loads behind the K reduction and
    // halves the loads in flight): with no visible key yet, mx == -inf and everything below is 0
    float const base = (mx == -INFINITY) ? 0.f : mx;
    float const corr = fast_exp2(m - base); // m == -inf -> 0 (l and acc are 0 then)
    l *= corr;
#pragma unroll
    for (int e = 0; e < 8; e++) {
      acc[e] *= corr;
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      float const p = fast_exp2(s[u] - base); // s == -inf -> 0
      l += p;
      float vf[8];
      unpack8<T>(vr[u], vf);
#pragma unroll
      for (int e = 0; e < 8; e++) {
        acc[e] = fmaf(p, vf[e], acc[e]);
      }
    }
    m = mx;
  }

  // merge the GROUPS independent (m, l, acc) streams of this block
  __shared__ float s_acc[GROUPS][DH];
  __shared__ float s_m[GROUPS], s_l[GROUPS];
  int const g = warp * KPW + sub;
#pragma unroll
  for (int e = 0; e < 8; e++) {
    s_acc[g][sl * 8 + e] = acc[e];
  }
  if (sl == 0) {
    s_m[g] = m;
    s_l[g] = l;
  }
  __syncthreads();
  __shared__ int s_last;
  if (threadIdx.x < DH) {
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < GROUPS; i++) {
      M = fmaxf(M, s_m[i]);
    }
    float L = 0.f, o = 0.f;
#pragma unroll
    for (int i = 0; i < GROUPS; i++) {
      float const w = (s_m[i] == -INFINITY) ? 0.f : fast_exp2(s_m[i] - M);
      L += s_l[i] * w;
      o += s_acc[i][threadIdx.x] * w;
    }
    if (nsplit == 1) {
      // reference normalisation: inv_sum = 1 / (exp_sum + 1e-6)  (inc_…cu:538)
      out[((size_t)t * nq + head) * DH + threadIdx.x] = DT<T>::from_f(o * __fdividef(1.f, L + 1.e-6f));
    } else {
      float *rec = partial + (((size_t)t * nq + head) * nsplit + split) * (DH + 2);
      rec[threadIdx.x] = o;
      if (threadIdx.x == 0) {
        rec[DH] = M;
        rec[DH + 1] = L;
      }
    }
  }
  if (nsplit > 1 && fuse_combine) {
    // last-arriving split of this (token, head) merges all records: no separate combine launch
    int const nvalid = min(nsplit, (kv_len + chunk - 1) / chunk);
    __syncthreads();
    if (threadIdx.x == 0) {
      // release/acquire at gpu scope; the bar.sync above makes the other threads' record stores
      // cumulative with this release, so no per-thread __threadfence is needed
      int prev;
      asm volatile("atom.acq_rel.gpu.global.add.s32 %0, [%1], 1;" : "=r"(prev) : "l"(split_counter + t * nq + head) : "memory");
      s_last = prev == nvalid - 1;
      if (s_last) {
        split_counter[t * nq + head] = 0; // self-reset for the next launch
      }
    }
    __syncthreads();
    if (s_last) {
      if (threadIdx.x < 32 * (DH / 32)) {
        combine_record<T, DH>(partial + ((size_t)t * nq + head) * nsplit * (DH + 2), nvalid,
                              out + ((size_t)t * nq + head) * DH, threadIdx.x, 32 * (DH / 32));
      }
    }
  }
}

// Combine the split-KV partials of one (token, head): exact log-sum-exp merge, then the reference's
// 1/(sum + 1e-6).  Shared by the decode and the multi-query kernels.
template <typename T, int DH>
__global__ void __launch_bounds__(DH)
combine_kernel(float const *__restrict__ partial, QueryDesc const *__restrict__ qdesc,
               T *__restrict__ out, int nq, int chunk, int nsplit) {
  int const head = blockIdx.x, t = blockIdx.y;
  int const kv_len = qdesc[t].kv_len;
  int const nvalid = min(nsplit, (kv_len + chunk - 1) / chunk);
  float const *rec = partial + ((size_t)t * nq + head) * nsplit * (DH + 2);
  float M = -INFINITY;
  for (int i = 0; i < nvalid; i++) {
    M = fmaxf(M, rec[i * (DH + 2) + DH]);
  }
  float L = 0.f, o = 0.f;
  for (int i = 0; i < nvalid; i++) {
    float const mi = rec[i * (DH + 2) + DH];
    float const w = (mi == -INFINITY) ? 0.f : fast_exp2(mi - M);
    L += rec[i * (DH + 2) + DH + 1] * w;
    o += rec[i * (DH + 2) + threadIdx.x] * w;
  }
  out[((size_t)t * nq + head) * DH + threadIdx.x] = DT<T>::from_f(o * __fdividef(1.f, L + 1.e-6f));
}

int launch_combine(ffb_attn_s *h, void *out, int T, int chunk, int nsplit, cudaStream_t s) {
  ffb_attn_desc const &d = h->d;
  dim3 grid(d.num_q_heads, T);
#define FFB_COMB(TT, DH) \
  combine_kernel<TT, DH><<<grid, DH, 0, s>>>(h->partial, h->qdesc, (TT *)out, d.num_q_heads, chunk, nsplit)
  if (d.dtype == FFB_DT_HALF) {
    if (d.head_dim == 128) FFB_COMB(__half, 128); else FFB_COMB(__half, 64);
  } else {
    if (d.head_dim == 128) FFB_COMB(__nv_bfloat16, 128); else FFB_COMB(__nv_bfloat16, 64);
  }
#undef FFB_COMB
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

// Split heuristic shared with the multi-query kernel: pick the largest chunk from {512,256,128,64}
// (keys per block) that still yields >= `target_ctas` blocks; cap by the partial-buffer capacity.
int choose_chunk(ffb_attn_s *h, long units, int kv_len_max, int granule, long target_ctas,
                 char const *env_name, int *nsplit_out) {
  int chunk = 64;
  int const cands[4] = {512, 256, 128, 64};
  for (int i = 0; i < 4; i++) {
    long n = units * ((kv_len_max + cands[i] - 1) / cands[i]);
    if (n >= target_ctas || i == 3) {
      chunk = cands[i];
      break;
    }
  }
  if (char const *e = ffb::env(env_name)) {
    int v = atoi(e);
    if (v > 0) {
      chunk = v;
    }
  }
  chunk = ((chunk + granule - 1) / granule) * granule;
  if (chunk >= kv_len_max) {
    *nsplit_out = 1;
    return chunk;
  }
  return (*nsplit_out = (kv_len_max + chunk - 1) / chunk), chunk;
}

int launch_decode_attn(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T,
                       int kv_len_max, cudaStream_t s) {
  (void)meta;
  ffb_attn_desc const &d = h->d;
  if (T <= 0) {
    return FFB_OK;
  }
  int const U = ffb::env("FFB_DECODE_U") ? atoi(ffb::env("FFB_DECODE_U")) : 8; // keys per warp-load group
  int const granule = kDecodeWarps * 8 * (32 / (d.head_dim / 8));
  int nsplit = 1;
  int chunk = choose_chunk(h, (long)T * d.num_q_heads, kv_len_max < 1 ? 1 : kv_len_max, granule,
                           (long)kNumSMs * 8, "FFB_DECODE_CHUNK", &nsplit);
  // respect the partial buffer
  size_t const rec = (size_t)(d.head_dim + 2) * sizeof(float);
  while (nsplit > 1 && (size_t)T * d.num_q_heads * nsplit * rec > h->partial_bytes) {
    chunk *= 2;
    nsplit = (kv_len_max + chunk - 1) / chunk;
  }
  float const scale = d.qk_prod_scaling ? 1.0f / sqrtf((float)d.head_dim) : 1.0f;
  float const scale_log2 = scale * 1.4426950408889634f;
  int const fuse = ffb::env("FFB_SEPARATE_COMBINE") ? 0 : 1;
  dim3 grid(nsplit, d.num_q_heads, T);
  if (h->profiling == 1) {
    cudaEventRecord(h->ev0, s);
  }
#define FFB_DEC_U(TT, DH, UU)                                                                             \
  decode_attn_kernel<TT, DH, UU><<<grid, kDecodeThreads, 0, s>>>(                                          \
      (TT const *)h->q_buf, (TT const *)h->k_cache, (TT const *)h->v_cache, h->qdesc, h->partial, h->split_counter, (TT *)out, \
      d.num_q_heads, d.num_kv_heads, h->cache_len, chunk, nsplit, scale_log2, fuse)
#define FFB_DEC(TT, DH)                     \
  do {                                      \
    if (U == 8) FFB_DEC_U(TT, DH, 8);       \
    else if (U == 2) FFB_DEC_U(TT, DH, 2);  \
    else FFB_DEC_U(TT, DH, 4);              \
  } while (0)
  if (d.dtype == FFB_DT_HALF) {
    if (d.head_dim == 128) FFB_DEC(__half, 128); else FFB_DEC(__half, 64);
  } else {
    if (d.head_dim == 128) FFB_DEC(__nv_bfloat16, 128); else FFB_DEC(__nv_bfloat16, 64);
  }
#undef FFB_DEC
#undef FFB_DEC_U
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
