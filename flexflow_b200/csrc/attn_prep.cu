// attn_prep.cu — fused RoPE + query staging + KV-cache append + per-query descriptors, and the
// commit of accepted tree tokens.  One pass over the [T][3][H] projection output replaces the
// reference's D2D copy + scaling_query_kernel + apply_rotary_embedding_hf + store_kv_cache /
// update_tree_branch_kv_cache_fused / spec_inc_store_kv_cache
// (src/ops/inc_multihead_self_attention.cu:35-61,650-738,944-962;
//  src/ops/tree_inc_multihead_self_attention.cu:433-478; src/ops/spec_inc_multihead_self_attention.cu:311-358).
// HBM-bound elementwise work: 128-bit loads/stores, sin/cos computed once per (token, pair)
// instead of once per (token, head, pair).
#include <cstdlib>

#include "attn_state.h"

namespace ffb {

template <typename T, int DH>
__global__ void __launch_bounds__(128)
prep_kernel(T const *__restrict__ qkv, T *__restrict__ q_buf, T *__restrict__ k_cache,
            T *__restrict__ v_cache, T *__restrict__ prev_kv, QueryDesc *__restrict__ qdesc,
            ffb_combined_batch_config_meta const *__restrict__ meta, int mode, int nq, int nkv,
            int cache_len, int max_requests, RopeParams rp, int scaling_query, float scaling_factor,
            int num_generation_tokens) {
  // the attention kernel that follows is launched as a programmatic dependent: let it start its set-up (and the
  // stream of K/V rows older than this batch) now; it waits for this grid before it touches anything written here
  asm volatile("griddepcontrol.wait;" ::: "memory"); // (no-op unless launched as a programmatic dependent)
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  int const t = blockIdx.x;
  int const H = nq * DH;
  int const G = nq / nkv;
  __shared__ float s_cos[DH / 2], s_sin[DH / 2];

  ffb_per_token_info const ti = meta->tokens_info[t];
  int const slot = ti.request_index;
  bool const slot_ok = slot >= 0 && slot < max_requests;
  int first_depth = 0, first_off = 0, ntok = 0;
  bool prompt_phase = false;
  if (slot_ok) {
    ffb_per_request_info const *ri = &meta->requestsInfo[slot];
    first_depth = ri->first_token_depth_in_request;
    first_off = ri->first_token_offset_in_batch;
    ntok = ri->num_tokens_in_batch;
    prompt_phase = ri->prompt_phase;
  }
  int const qi = t - first_off;
  int cache_pos, kv_len, ntcs = 0, bit = -1;
  if (mode == FFB_INC_DECODING_MODE) {
    // store_kv_cache: cache[req][abs_depth]; generation kernel tlength (inc_…cu:420-422);
    // prompt path causal rule (inc_…cu:228-245)
    cache_pos = ti.abs_depth_in_request;
    kv_len = prompt_phase ? first_depth + qi + 1 : first_depth + ntok;
  } else if (mode == FFB_TREE_VERIFY_MODE) {
    // update_tree_branch_kv_cache_fused: slot = position in batch (tree_…cu:464-477);
    // mask rule tree_…cu:165-169
    cache_pos = first_depth + qi;
    int const tlength = first_depth + ntok;
    if (prompt_phase) {
      kv_len = min(tlength, first_depth + qi + 1);
    } else {
      kv_len = tlength;
      ntcs = slot_ok ? meta->causalMask[slot].non_tree_cache_size : 0;
      bit = qi;
    }
  } else {
    // spec_inc_store_kv_cache (spec_…cu:343-346); generation kernel (spec_…cu:87-89,143-144,165-167)
    int prompt_size = 0, tree_size = 0, this_layer = 0, nt = 0;
    if (slot_ok) {
      ffb_bit_mask const *bm = &meta->causalMask[slot];
      prompt_size = bm->prompt_size; tree_size = bm->tree_size;
      this_layer = bm->this_layer_size; nt = bm->non_tree_cache_size;
    }
    cache_pos = prompt_size + nt + tree_size - 1 - this_layer + qi;
    if (t < num_generation_tokens) {
      int const branches = slot_ok ? meta->beamRequestsInfo[slot].sub_request_num : 0;
      kv_len = nt + tree_size + prompt_size - 1;
      ntcs = nt;
      bit = prompt_size + tree_size - 1 - branches + qi;
    } else {
      kv_len = first_depth + qi + 1;
    }
  }
  bool const pos_ok = slot_ok && cache_pos >= 0 && cache_pos < cache_len;
  if (threadIdx.x < 32 && blockIdx.y == 0) {
    // warp 0 transposes the query's column of the request's tree mask: lane i holds mask[i] and mask[i + 32]
    unsigned long long vis = 0ULL;
    if (bit >= 0 && bit < 64 && slot_ok) { // (uniform across the warp)
      unsigned long long const m0 = meta->causalMask[slot].mask[threadIdx.x];
      unsigned long long const m1 = meta->causalMask[slot].mask[threadIdx.x + 32];
      uint32_t const lo = __ballot_sync(0xffffffffu, (m0 >> bit) & 1ULL);
      uint32_t const hi = __ballot_sync(0xffffffffu, (m1 >> bit) & 1ULL);
      vis = ((unsigned long long)hi << 32) | lo;
    }
    if (threadIdx.x == 0) {
      QueryDesc qd;
      qd.slot = slot_ok ? slot : 0;
      qd.kv_len = slot_ok ? max(0, min(kv_len, cache_len)) : 0;
      qd.ntcs = ntcs;
      qd.bit = (bit >= 0 && bit < 64) ? bit : (bit < 0 ? -1 : 64); // 64: nothing beyond ntcs visible
      qd.vis = vis;
      qdesc[t] = qd;
    }
  }

  // sin/cos of this token's position, once per complex pair (reference: per head and pair)
  for (int i = threadIdx.x; i < DH / 2; i += blockDim.x) {
    float c = 1.f, s = 0.f;
    if (rp.apply) {
      size_t pos = (size_t)ti.abs_depth_in_request;
      // theta_i as in apply_rotary_embedding_hf (inc_…cu:708-709): float pow, double divide
      float freq = pos * (1.0 / pow(rp.theta, (float)2 * i / DH));
      if (rp.llama3) { // inc_…cu:711-730
        float const pi = 3.14159265358979323846f;
        float wavelen = 2 * pi / freq;
        float low_freq_wavelen = rp.orig_max_pos / rp.low_ff;
        float high_freq_wavelen = rp.orig_max_pos / rp.high_ff;
        if (wavelen < high_freq_wavelen) {
        } else if (wavelen > low_freq_wavelen) {
          freq = freq / rp.factor;
        } else {
          float smooth = (rp.orig_max_pos / wavelen - rp.low_ff) / (rp.high_ff - rp.low_ff);
          freq = ((1 - smooth) * freq / rp.factor + smooth * freq);
        }
      }
      c = cosf(freq);
      s = sinf(freq);
    }
    s_cos[i] = c;
    s_sin[i] = s;
  }
  __syncthreads();

  constexpr int CH = DH / 16; // 16-byte chunks per half head
  T const *row = qkv + (size_t)t * 3 * H;
  // rotate Q heads (-> q_buf) and the distinct K heads (-> cache, prev_kv)
  // work items of this token (rotate-Q/K chunks, then V chunks) are spread over gridDim.y blocks
  int const n_rot = (nq + nkv) * CH;
  int const n_v = nkv * (DH / 8);
  int const tstride = blockDim.x * gridDim.y;
  for (int item = blockIdx.y * blockDim.x + threadIdx.x; item < n_rot + n_v; item += tstride) {
    if (item < n_rot) {
      int const head = item / CH, c = item % CH;
      bool const is_q = head < nq;
      int const src_head = is_q ? head : (head - nq) * G;
      T const *src = row + (is_q ? 0 : H) + src_head * DH + c * 8;
      uint4 ux = *reinterpret_cast<uint4 const *>(src);
      uint4 uy = *reinterpret_cast<uint4 const *>(src + DH / 2);
      float x[8], y[8], ox[8], oy[8];
      unpack8<T>(ux, x);
      unpack8<T>(uy, y);
#pragma unroll
      for (int e = 0; e < 8; e++) {
        float xv = x[e], yv = y[e];
        if (is_q && scaling_query) { // scaling_query_kernel stores DT (inc_…cu:650-662)
          xv = DT<T>::to_f(DT<T>::from_f(xv * scaling_factor));
          yv = DT<T>::to_f(DT<T>::from_f(yv * scaling_factor));
        }
        float cs = s_cos[c * 8 + e], sn = s_sin[c * 8 + e];
        ox[e] = xv * cs - yv * sn; // cuCmulf
        oy[e] = xv * sn + yv * cs;
      }
      uint4 px = pack8<T>(ox), py = pack8<T>(oy);
      if (is_q) {
        T *dst = q_buf + ((size_t)t * nq + head) * DH + c * 8;
        *reinterpret_cast<uint4 *>(dst) = px;
        *reinterpret_cast<uint4 *>(dst + DH / 2) = py;
      } else {
        int const j = head - nq;
        if (pos_ok) {
          T *dst = k_cache + (((size_t)slot * nkv + j) * cache_len + cache_pos) * DH + c * 8;
          *reinterpret_cast<uint4 *>(dst) = px;
          *reinterpret_cast<uint4 *>(dst + DH / 2) = py;
        }
        if (prev_kv) {
          T *dst = prev_kv + (((size_t)t * 2 + 0) * nkv + j) * DH + c * 8;
          *reinterpret_cast<uint4 *>(dst) = px;
          *reinterpret_cast<uint4 *>(dst + DH / 2) = py;
        }
      }
    } else { // V heads: straight copy
      int const w = item - n_rot;
      int const j = w / (DH / 8), c = w % (DH / 8);
      uint4 u = *reinterpret_cast<uint4 const *>(row + 2 * H + j * G * DH + c * 8);
      if (pos_ok) {
        *reinterpret_cast<uint4 *>(v_cache + (((size_t)slot * nkv + j) * cache_len + cache_pos) * DH + c * 8) = u;
      }
      if (prev_kv) {
        *reinterpret_cast<uint4 *>(prev_kv + (((size_t)t * 2 + 1) * nkv + j) * DH + c * 8) = u;
      }
    }
  }
}

// commit_tokens_kernel (tree_…cu:335-370): cache[request_index][token_depth] <- K,V of token
// `token_index` of the PREVIOUS verify batch.
template <typename T>
__global__ void commit_kernel(T const *__restrict__ prev_kv, T *__restrict__ k_cache,
                              T *__restrict__ v_cache,
                              ffb_combined_batch_config_meta const *__restrict__ meta, int nkv, int dh,
                              int cache_len, int max_requests, int prev_tokens) {
  ffb_committed_token_info const ci = meta->committed_tokens[blockIdx.x];
  if (ci.token_index < 0 || ci.token_index >= prev_tokens || ci.request_index < 0 ||
      ci.request_index >= max_requests || ci.token_depth < 0 || ci.token_depth >= cache_len) {
    return; // the reference asserts (tree_…cu:353)
  }
  int const per = nkv * dh / 8; // uint4 per K (or V) row set
  for (int w = threadIdx.x; w < 2 * per; w += blockDim.x) {
    int const kv = w / per, r = w % per;
    int const j = r / (dh / 8), c = r % (dh / 8);
    uint4 u = *reinterpret_cast<uint4 const *>(prev_kv + (((size_t)ci.token_index * 2 + kv) * nkv + j) * dh + c * 8);
    T *cache = kv == 0 ? k_cache : v_cache;
    *reinterpret_cast<uint4 *>(cache + (((size_t)ci.request_index * nkv + j) * cache_len + ci.token_depth) * dh + c * 8) = u;
  }
}

// test/diagnostic access: [n][nkv][dh] <-> cache rows
template <typename T>
__global__ void kv_io_kernel(T *k_cache, T *v_cache, T *k, T *v, int request, int pos0, int n, int nkv,
                             int dh, int cache_len, bool write) {
  size_t total = (size_t)n * nkv * dh;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    int d = i % dh, j = (i / dh) % nkv, p = i / ((size_t)dh * nkv);
    size_t ci = (((size_t)request * nkv + j) * cache_len + pos0 + p) * dh + d;
    if (write) {
      if (k) k_cache[ci] = k[i];
      if (v) v_cache[ci] = v[i];
    } else {
      if (k) k[i] = k_cache[ci];
      if (v) v[i] = v_cache[ci];
    }
  }
}

template <typename T>
static int launch_prep_t(ffb_attn_s *h, int mode, ffb_combined_batch_config_meta const *meta, void const *qkv,
                         int T_, int ngen, cudaStream_t s) {
  ffb_attn_desc const &d = h->d;
  RopeParams rp{d.apply_rotary_embedding, d.rope_theta, d.rope_llama3, d.rope_factor,
                d.rope_low_freq_factor, d.rope_high_freq_factor, d.rope_original_max_position_embeddings};
  T *prev = mode == FFB_TREE_VERIFY_MODE ? (T *)h->prev_kv : nullptr;
  // enough blocks to cover ~2 waves of SMs, at most one item per thread
  int const items = (d.num_q_heads + d.num_kv_heads) * (d.head_dim / 16) + d.num_kv_heads * (d.head_dim / 8);
  int pb = (items + 127) / 128;
  while (pb > 1 && (long)T_ * pb > 4L * kNumSMs) {
    pb = (pb + 1) / 2;
  }
  // experiment hook: FFB_PREP_CARVEOUT=<percent> asks for that shared-memory carve-out, so that a small kernel between
  // two 200 KB-shared-memory kernels does not flip the SMs' L1 / shared split twice
  static int const carve = ffb::env("FFB_PREP_CARVEOUT") ? atoi(ffb::env("FFB_PREP_CARVEOUT")) : -1;
  // FFB_PREP_PDL=1 launches this kernel as a programmatic dependent of the projection GEMM before it.  OFF by default:
  // measured at configs[3] the attention kernel (itself a dependent of this one, triggered at its top) then becomes
  // resident first and its spinning warps share the SMs with this grid: prep + attention 42.8 -> 50.9 us.
  static bool const prep_pdl = ffb::env("FFB_PREP_PDL") && atoi(ffb::env("FFB_PREP_PDL")) != 0;
#define FFB_PREP(DH)                                                                                   \
  if (carve >= 0) {                                                                                    \
    static bool set = false;                                                                           \
    if (!set) {                                                                                        \
      cudaFuncSetAttribute(prep_kernel<T, DH>, cudaFuncAttributePreferredSharedMemoryCarveout, carve); \
      set = true;                                                                                      \
    }                                                                                                  \
  }                                                                                                    \
  {                                                                                                    \
    cudaLaunchConfig_t cfg = {};                                                                       \
    cfg.gridDim = dim3(T_, pb);                                                                        \
    cfg.blockDim = dim3(128);                                                                          \
    cfg.stream = s;                                                                                    \
    cudaLaunchAttribute attr[1];                                                                       \
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                   \
    attr[0].val.programmaticStreamSerializationAllowed = 1;                                            \
    cfg.attrs = attr;                                                                                  \
    cfg.numAttrs = prep_pdl ? 1 : 0;                                                                   \
    cudaLaunchKernelEx(&cfg, prep_kernel<T, DH>, (T const *)qkv, (T *)h->q_buf, (T *)h->k_cache, (T *)h->v_cache, \
                       (T *)prev, (QueryDesc *)h->qdesc, meta, mode, (int)d.num_q_heads, (int)d.num_kv_heads, \
                       (int)h->cache_len, (int)d.max_requests, rp, (int)d.scaling_query, (float)d.scaling_factor, \
                       (int)ngen);                                                                     \
  }
  if (d.head_dim == 128) {
    FFB_PREP(128);
  } else if (d.head_dim == 64) {
    FFB_PREP(64);
  } else {
    set_error("unsupported head_dim %d", d.head_dim);
    return FFB_ERR_INVALID;
  }
#undef FFB_PREP
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

int launch_prep(ffb_attn_s *h, int mode, ffb_combined_batch_config_meta const *meta, void const *qkv, int T,
                int num_generation_tokens, int speculative_request_num, cudaStream_t s) {
  (void)speculative_request_num;
  if (T <= 0) {
    return FFB_OK;
  }
  return h->d.dtype == FFB_DT_HALF ? launch_prep_t<__half>(h, mode, meta, qkv, T, num_generation_tokens, s)
                                   : launch_prep_t<__nv_bfloat16>(h, mode, meta, qkv, T, num_generation_tokens, s);
}

int launch_commit(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, int n_commit, cudaStream_t s) {
  if (n_commit <= 0) {
    return FFB_OK;
  }
  ffb_attn_desc const &d = h->d;
  // element type is irrelevant for a copy: use __half for both 16-bit types
  commit_kernel<__half><<<n_commit, 128, 0, s>>>((__half const *)h->prev_kv, (__half *)h->k_cache,
                                                 (__half *)h->v_cache, meta, d.num_kv_heads, d.head_dim,
                                                 h->cache_len, d.max_requests, h->num_active_infr_tokens);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

int launch_kv_io(ffb_attn_s *h, int request, int pos0, int n, void *k, void *v, bool write, cudaStream_t s) {
  if (n <= 0) {
    return FFB_OK;
  }
  ffb_attn_desc const &d = h->d;
  size_t total = (size_t)n * d.num_kv_heads * d.head_dim;
  int blocks = (int)((total + 255) / 256);
  blocks = blocks > 148 * 8 ? 148 * 8 : blocks;
  kv_io_kernel<__half><<<blocks, 256, 0, s>>>((__half *)h->k_cache, (__half *)h->v_cache, (__half *)k,
                                              (__half *)v, request, pos0, n, d.num_kv_heads, d.head_dim,
                                              h->cache_len, write);
  FFB_LAUNCH_CHECK();
  return FFB_OK;
}

} // namespace ffb
