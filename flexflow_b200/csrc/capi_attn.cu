// capi_attn.cu — C ABI for the three attention operators (include/ffb200.h).
// Host-side orchestration mirrors Kernels::IncMultiHeadAttention::inference_kernel
// (src/ops/inc_multihead_self_attention.cu:944-985), TreeIncMultiHeadAttention::inference_kernel
// (src/ops/tree_inc_multihead_self_attention.cu:575-626) and SpecIncMultiHeadSelfAttention::inference_kernel
// (src/ops/spec_inc_multihead_self_attention.cu:681-724), minus the two D2D staging copies.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <new>

#include <mutex>
#include <vector>

#include "attn_state.h"

namespace ffb {
static thread_local char g_err[512] = "";
void set_error(char const *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
namespace {
struct EnvSlot {
  std::atomic<char const *> key{nullptr};
  std::atomic<char const *> val{nullptr};
  std::atomic<int> gen{-1};
};
constexpr int kEnvSlots = 256;
EnvSlot g_env[kEnvSlots];
std::atomic<int> g_env_gen{0};
} // namespace

char const *env(char const *name) {
  int const gen = g_env_gen.load(std::memory_order_acquire);
  size_t const h0 = ((uintptr_t)name >> 2) % kEnvSlots;
  for (int i = 0; i < kEnvSlots; i++) {
    EnvSlot &sl = g_env[(h0 + i) % kEnvSlots];
    char const *k = sl.key.load(std::memory_order_acquire);
    if (k == nullptr) {
      char const *expected = nullptr;
      if (!sl.key.compare_exchange_strong(expected, name, std::memory_order_acq_rel) && expected != name) {
        continue; // another thread claimed the slot for a different name
      }
      k = name;
    }
    if (k == name) {
      if (sl.gen.load(std::memory_order_acquire) != gen) {
        char const *v = getenv(name);
        sl.val.store(v ? strdup(v) : nullptr, std::memory_order_release); // (a few bytes per reload are never freed)
        sl.gen.store(gen, std::memory_order_release);
      }
      return sl.val.load(std::memory_order_acquire);
    }
  }
  return getenv(name); // table full: uncached
}
} // namespace ffb

using namespace ffb;

extern "C" {

void ffb_reload_env(void) { g_env_gen.fetch_add(1, std::memory_order_acq_rel); }

char const *ffb_last_error(void) { return g_err; }

char const *ffb_version(void) {
  return "ffb200 0.1 (sm_100a, CUDA " FFB_STR(CUDART_VERSION) ")";
}

int ffb_check_device(void) {
  // called at the top of the stand-alone row operators: remember which devices passed (cudaGetDeviceProperties is slow)
  static std::atomic<unsigned long long> ok_mask{0};
  int dev = 0;
  FFB_CUDA_TRY(cudaGetDevice(&dev));
  if (dev >= 0 && dev < 64 && ((ok_mask.load(std::memory_order_acquire) >> dev) & 1ULL)) {
    return FFB_OK;
  }
  int major = 0, minor = 0;
  FFB_CUDA_TRY(cudaDeviceGetAttribute(&major, cudaDevAttrComputeCapabilityMajor, dev));
  FFB_CUDA_TRY(cudaDeviceGetAttribute(&minor, cudaDevAttrComputeCapabilityMinor, dev));
  if (major != 10) {
    set_error("device %d is sm_%d%d; this library only contains sm_100a code and has no fallback", dev, major, minor);
    return FFB_ERR_UNSUPPORTED;
  }
  if (dev >= 0 && dev < 64) {
    ok_mask.fetch_or(1ULL << dev, std::memory_order_release);
  }
  return FFB_OK;
}

// Multi-query attention: the tcgen05 / TMA kernel (attn_tc.cu) by default; FFB_ATTN_TC=0 selects the mma.sync
// kernel (attn_multi.cu), kept as the in-tree comparison point.
static int launch_multi_dispatch(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T,
                                 int kv_len_max, int q_len_max, AtcReq const *reqs, int nreq, cudaStream_t s) {
  char const *e = ffb::env("FFB_ATTN_TC");
  bool const want_tc = !(e && e[0] == '0');
  if (want_tc && attn_tc_supported(h)) {
    int const rc = launch_attn_tc(h, meta, out, T, reqs, nreq, s);
    if (rc != -1) {
      return rc;
    } // -1: the batch does not fit the schedule table
  }
  return launch_multi_attn(h, meta, out, T, kv_len_max, q_len_max, s);
}

int ffb_attn_create(ffb_attn_desc const *desc, ffb_attn_t *out) {
  FFB_REQUIRE(desc && out, "null argument");
  ffb_attn_desc const &d = *desc;
  FFB_REQUIRE(d.mode == FFB_INC_DECODING_MODE || d.mode == FFB_BEAM_SEARCH_MODE || d.mode == FFB_TREE_VERIFY_MODE,
              "unknown inference mode %d", d.mode);
  FFB_REQUIRE(d.dtype == FFB_DT_HALF || d.dtype == FFB_DT_BF16, "dtype %d not supported (half/bf16 only)", d.dtype);
  FFB_REQUIRE(d.head_dim == 64 || d.head_dim == 128, "head_dim %d not supported (64 or 128)", d.head_dim);
  FFB_REQUIRE(d.num_q_heads > 0 && d.num_kv_heads > 0 && d.num_q_heads % d.num_kv_heads == 0,
              "num_q_heads %d must be a positive multiple of num_kv_heads %d", d.num_q_heads, d.num_kv_heads);
  FFB_REQUIRE(d.max_requests > 0 && d.max_requests <= FFB_MAX_NUM_REQUESTS, "max_requests %d out of range", d.max_requests);
  FFB_REQUIRE(d.max_tokens_per_batch > 0 && d.max_tokens_per_batch <= FFB_MAX_NUM_TOKENS, "max_tokens_per_batch %d out of range", d.max_tokens_per_batch);
  FFB_REQUIRE(d.max_sequence_length > 0 && d.max_spec_tree_tokens >= 0 && d.max_spec_tree_tokens <= FFB_MAX_SPEC_TREE_TOKEN_NUM,
              "bad sequence limits");
  if (d.position_bias) {
    set_error("position_bias (ALiBi, MPT only) is outside the LLaMA hot path");
    return FFB_ERR_UNSUPPORTED;
  }
  int rc = ffb_check_device();
  if (rc != FFB_OK) {
    return rc;
  }
  ffb_attn_s *h = new (std::nothrow) ffb_attn_s();
  FFB_REQUIRE(h, "out of host memory");
  h->d = d;
  h->G = d.num_q_heads / d.num_kv_heads;
  h->elt = 2;
  // inc_…cu:1691-1716: beam/tree modes add max_spec_tree_token_num rows per request
  h->cache_len = d.max_sequence_length + (d.mode == FFB_INC_DECODING_MODE ? 0 : d.max_spec_tree_tokens);
  auto align = [](size_t x) { return (x + 255) & ~(size_t)255; };
  size_t const cache_b = align((size_t)d.max_requests * d.num_kv_heads * h->cache_len * d.head_dim * h->elt);
  size_t const q_b = align((size_t)d.max_tokens_per_batch * d.num_q_heads * d.head_dim * h->elt);
  size_t const prev_b = d.mode == FFB_TREE_VERIFY_MODE
                            ? align((size_t)d.max_tokens_per_batch * 2 * d.num_kv_heads * d.head_dim * h->elt) : 0;
  size_t const qd_b = align((size_t)d.max_tokens_per_batch * sizeof(QueryDesc));
  size_t const cnt_b = align((size_t)d.max_tokens_per_batch * d.num_q_heads * sizeof(int));
  size_t const rec = (size_t)(d.head_dim + 2) * sizeof(float);
  size_t part = (size_t)d.max_tokens_per_batch * d.num_q_heads * 8 * rec;
  size_t const part_min = (size_t)d.num_q_heads * ((h->cache_len + 63) / 64) * rec * 2;
  if (part > ((size_t)96 << 20)) {
    part = (size_t)96 << 20;
  }
  // the tcgen05 multi-query kernel keeps two split-record slots of [Dh + 2][128] floats per CTA (attn_tc.cu)
  size_t const part_tc = (size_t)2 * kNumSMs * (d.head_dim + 2) * 128 * sizeof(float);
  if (part < part_tc) {
    part = part_tc;
  }
  if (part < part_min) {
    part = part_min;
  }
  h->partial_bytes = align(part);
  h->bytes = 2 * cache_b + q_b + prev_b + qd_b + cnt_b + h->partial_bytes;
  cudaError_t e = cudaMalloc(&h->base, h->bytes);
  if (e != cudaSuccess) {
    set_error("cudaMalloc(%zu bytes) failed: %s", h->bytes, cudaGetErrorString(e));
    delete h;
    return FFB_ERR_CUDA;
  }
  uint8_t *p = (uint8_t *)h->base;
  h->k_cache = p; p += cache_b;
  h->v_cache = p; p += cache_b;
  h->q_buf = p; p += q_b;
  h->prev_kv = prev_b ? p : nullptr; p += prev_b;
  h->qdesc = (QueryDesc *)p; p += qd_b;
  h->split_counter = (int *)p; p += cnt_b;
  h->partial = (float *)p;
  e = cudaMemset(h->base, 0, h->bytes);
  if (e != cudaSuccess) {
    set_error("cudaMemset failed: %s", cudaGetErrorString(e));
    cudaFree(h->base);
    delete h;
    return FFB_ERR_CUDA;
  }
  rc = attn_tc_init(h);
  if (rc != FFB_OK) {
    cudaFree(h->base);
    delete h;
    return rc;
  }
  *out = h;
  return FFB_OK;
}

int ffb_attn_destroy(ffb_attn_t h) {
  if (!h) {
    return FFB_OK;
  }
  if (h->ev0) {
    cudaEventDestroy(h->ev0);
    cudaEventDestroy(h->ev1);
  }
  cudaError_t e = cudaFree(h->base);
  delete h;
  if (e != cudaSuccess) {
    set_error("cudaFree failed: %s", cudaGetErrorString(e));
    return FFB_ERR_CUDA;
  }
  return FFB_OK;
}

size_t ffb_attn_device_bytes(ffb_attn_t h) { return h ? h->bytes : 0; }

int ffb_attn_set_profiling(ffb_attn_t h, int enable) {
  FFB_REQUIRE(h, "null argument");
  if (enable && !h->ev0) {
    FFB_CUDA_TRY(cudaEventCreate(&h->ev0));
    FFB_CUDA_TRY(cudaEventCreate(&h->ev1));
  }
  h->profiling = enable == 2 ? 2 : (enable != 0);
  return FFB_OK;
}

int ffb_attn_last_kernel_ms(ffb_attn_t h, float *ms) {
  FFB_REQUIRE(h && ms && h->ev0, "profiling was never enabled on this handle");
  FFB_CUDA_TRY(cudaEventSynchronize(h->ev1));
  FFB_CUDA_TRY(cudaEventElapsedTime(ms, h->ev0, h->ev1));
  return FFB_OK;
}

} // extern "C"

namespace {
// fence events of the staging blocks handed to ffb_load_batch_config, keyed by (block, device)
struct StagingFence {
  void *staging;
  int device;
  cudaEvent_t ev;
};
std::mutex g_fence_mu;
std::vector<StagingFence> g_fences;

// Waits until the previous upload from `staging` on the current device has been executed and returns the event to
// record after the next one.
int staging_fence(void *staging, cudaEvent_t *out) {
  int dev = 0;
  FFB_CUDA_TRY(cudaGetDevice(&dev));
  cudaEvent_t ev = nullptr;
  {
    std::lock_guard<std::mutex> lock(g_fence_mu);
    for (StagingFence const &f : g_fences) {
      if (f.staging == staging && f.device == dev) {
        ev = f.ev;
        break;
      }
    }
    if (!ev) {
      FFB_CUDA_TRY(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
      g_fences.push_back(StagingFence{staging, dev, ev});
      *out = ev;
      return FFB_OK; // first use of this block: nothing in flight
    }
  }
  FFB_CUDA_TRY(cudaEventSynchronize(ev));
  *out = ev;
  return FFB_OK;
}
} // namespace

extern "C" {

int ffb_load_batch_config(void const *bc_host, int mode, void *staging, ffb_combined_batch_config_meta *meta_dev,
                          ffb_stream_t stream) {
  FFB_REQUIRE(bc_host && staging && meta_dev, "null argument");
  ffb_batch_config const *bc = (ffb_batch_config const *)bc_host;
  ffb_combined_batch_config_meta *st = (ffb_combined_batch_config_meta *)staging;
  FFB_REQUIRE(bc->num_tokens >= 0 && bc->num_tokens <= FFB_MAX_NUM_TOKENS, "num_tokens %d out of range", bc->num_tokens);
  // The copy below is truly asynchronous (pinned source), unlike the reference's pageable copies, which the driver
  // stages before returning: a host that runs ahead of the stream must not repack a staging block whose previous
  // copy has not run yet.  Each (staging block, device) pair therefore carries a fence event, recorded after its copy
  // and waited for here before the block is written again.  (Skipped under stream capture: events recorded into a
  // graph cannot be waited for from the host; a captured upload replays its capture-time bytes anyway.)
  cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
  cudaStreamIsCapturing((cudaStream_t)stream, &cap);
  bool const capturing = cap != cudaStreamCaptureStatusNone;
  cudaEvent_t fence = nullptr;
  if (!capturing) {
    int rc = staging_fence(staging, &fence);
    if (rc != FFB_OK) {
      return rc;
    }
  }
  // request_manager.cu:98-164 copies the same members one cudaMemcpyAsync at a time
  memcpy(st->tokens_info, bc->tokensInfo, sizeof(ffb_per_token_info) * (size_t)bc->num_tokens);
  memcpy(st->requestsInfo, bc->requestsInfo, sizeof(st->requestsInfo));
  memcpy(st->causalMask, bc->causalMask, sizeof(st->causalMask));
  memcpy(st->request_completed, bc->request_completed, sizeof(st->request_completed));
  size_t bytes = offsetof(ffb_combined_batch_config_meta, beamTokenInfo);
  if (mode == FFB_BEAM_SEARCH_MODE) {
    ffb_beam_search_batch_config const *b = (ffb_beam_search_batch_config const *)bc_host;
    memcpy(st->beamTokenInfo, b->beamTokenInfo, sizeof(st->beamTokenInfo));
    memcpy(st->beamRequestsInfo, b->beamRequestsInfo, sizeof(st->beamRequestsInfo));
    bytes = offsetof(ffb_combined_batch_config_meta, committed_tokens);
  } else if (mode == FFB_TREE_VERIFY_MODE) {
    ffb_tree_verify_batch_config const *t = (ffb_tree_verify_batch_config const *)bc_host;
    FFB_REQUIRE(t->num_tokens_to_commit >= 0 && t->num_tokens_to_commit <= FFB_MAX_NUM_TOKENS, "num_tokens_to_commit out of range");
    memcpy(st->committed_tokens, t->committed_tokens, sizeof(ffb_committed_token_info) * (size_t)t->num_tokens_to_commit);
    bytes = sizeof(ffb_combined_batch_config_meta);
  } else {
    FFB_REQUIRE(mode == FFB_INC_DECODING_MODE, "unknown inference mode %d", mode);
  }
  FFB_CUDA_TRY(cudaMemcpyAsync(meta_dev, st, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  if (fence && !capturing) {
    FFB_CUDA_TRY(cudaEventRecord(fence, (cudaStream_t)stream));
  }
  return FFB_OK;
}

} // extern "C"

namespace {

struct BatchScan {
  int kv_len_max = 0, q_len_max = 0;
  bool any_multi = false; // some request contributes more than one query row per head
  int nreq = 0;
  AtcReq reqs[FFB_MAX_NUM_REQUESTS];
  void add(int slot, ffb_per_request_info const &ri, int qlen, int tl, int ntcs, bool causal) {
    AtcReq &r = reqs[nreq++];
    r.slot = slot; r.first_tok = ri.first_token_offset_in_batch; r.qlen = qlen; r.tl = tl; r.ntcs = ntcs;
    r.causal = causal ? 1 : 0; r.first_depth = ri.first_token_depth_in_request;
  }
};

int check_common(ffb_attn_t h, ffb_batch_config const *bc, void const *meta, void const *qkv, void *out, int mode) {
  FFB_REQUIRE(h && bc && meta && qkv && out, "null argument");
  FFB_REQUIRE(h->d.mode == mode, "handle was created for mode %d, called as %d", h->d.mode, mode);
  FFB_REQUIRE(bc->num_tokens >= 0 && bc->num_tokens <= h->d.max_tokens_per_batch,
              "num_tokens %d exceeds max_tokens_per_batch %d", bc->num_tokens, h->d.max_tokens_per_batch);
  for (int i = 0; i < h->d.max_requests; i++) {
    if (!bc->request_completed[i] && bc->requestsInfo[i].peft_bwd) {
      set_error("request %d asks for PEFT backward: training is outside the hot path", i);
      return FFB_ERR_UNSUPPORTED;
    }
  }
  return FFB_OK;
}

int forced_kernel() { // FFB_FORCE_ATTN=decode|multi (tests / profiling)
  char const *e = ffb::env("FFB_FORCE_ATTN");
  if (!e) {
    return 0;
  }
  return strcmp(e, "decode") == 0 ? 1 : (strcmp(e, "multi") == 0 ? 2 : 0);
}

} // namespace

extern "C" {

int ffb_inc_attn_forward(ffb_attn_t h, ffb_batch_config const *bc, ffb_combined_batch_config_meta const *meta,
                         int shard_id, void const *qkv, void *out, ffb_stream_t stream) {
  (void)shard_id; // only used for ALiBi head offsets in the reference (inc_…cu:221)
  int rc = check_common(h, bc, meta, qkv, out, FFB_INC_DECODING_MODE);
  if (rc != FFB_OK) {
    return rc;
  }
  int const T = bc->num_tokens;
  if (T == 0) {
    return FFB_OK;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BatchScan sc;
  for (int i = 0; i < h->d.max_requests; i++) {
    if (bc->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &ri = bc->requestsInfo[i];
    int const tl = ri.first_token_depth_in_request + ri.num_tokens_in_batch;
    FFB_REQUIRE(tl <= h->cache_len, "request %d: %d tokens exceed the cache length %d", i, tl, h->cache_len);
    sc.kv_len_max = tl > sc.kv_len_max ? tl : sc.kv_len_max;
    sc.q_len_max = ri.num_tokens_in_batch > sc.q_len_max ? ri.num_tokens_in_batch : sc.q_len_max;
    sc.any_multi |= ri.num_tokens_in_batch > 1;
    sc.add(i, ri, ri.num_tokens_in_batch, tl, bc->causalMask[i].non_tree_cache_size, ri.prompt_phase);
  }
  h->num_generation_tokens_cur = bc->num_generation_tokens;
  rc = h->profiling == 2 ? FFB_OK : launch_prep(h, FFB_INC_DECODING_MODE, meta, qkv, T, bc->num_generation_tokens, 0, s);
  if (rc != FFB_OK) {
    return rc;
  }
  int const force = forced_kernel();
  bool const use_decode = force == 1 || (force == 0 && !sc.any_multi && h->G == 1);
  return use_decode ? launch_decode_attn(h, meta, out, T, sc.kv_len_max, s)
                    : launch_multi_dispatch(h, meta, out, T, sc.kv_len_max, sc.q_len_max, sc.reqs, sc.nreq, s);
}

int ffb_tree_attn_forward(ffb_attn_t h, ffb_tree_verify_batch_config const *tbc,
                          ffb_combined_batch_config_meta const *meta, int shard_id, void const *qkv, void *out,
                          ffb_stream_t stream) {
  (void)shard_id;
  ffb_batch_config const *bc = (ffb_batch_config const *)tbc;
  int rc = check_common(h, bc, meta, qkv, out, FFB_TREE_VERIFY_MODE);
  if (rc != FFB_OK) {
    return rc;
  }
  cudaStream_t s = (cudaStream_t)stream;
  FFB_REQUIRE(tbc->num_tokens_to_commit >= 0 && tbc->num_tokens_to_commit <= FFB_MAX_NUM_TOKENS,
              "num_tokens_to_commit %d out of range", tbc->num_tokens_to_commit);
  for (int c = 0; c < tbc->num_tokens_to_commit; c++) { // the reference asserts on device (tree_…cu:353)
    if (tbc->committed_tokens[c].token_index >= h->num_active_infr_tokens || tbc->committed_tokens[c].token_index < 0) {
      set_error("committed token %d refers to index %d but the previous batch had %d tokens", c,
                tbc->committed_tokens[c].token_index, h->num_active_infr_tokens);
      return FFB_ERR_STATE;
    }
    FFB_REQUIRE(tbc->committed_tokens[c].token_depth >= 0 && tbc->committed_tokens[c].token_depth < h->cache_len &&
                    tbc->committed_tokens[c].request_index >= 0 && tbc->committed_tokens[c].request_index < h->d.max_requests,
                "committed token %d out of range", c);
  }
  rc = launch_commit(h, meta, tbc->num_tokens_to_commit, s);
  if (rc != FFB_OK) {
    return rc;
  }
  int const T = bc->num_tokens;
  h->num_active_infr_tokens = T; // tree_…cu:594
  if (T == 0) {
    return FFB_OK;
  }
  BatchScan sc;
  for (int i = 0; i < h->d.max_requests; i++) {
    if (bc->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &ri = bc->requestsInfo[i];
    int const tl = ri.first_token_depth_in_request + ri.num_tokens_in_batch;
    FFB_REQUIRE(tl <= h->cache_len, "request %d: %d tokens exceed the cache length %d", i, tl, h->cache_len);
    sc.kv_len_max = tl > sc.kv_len_max ? tl : sc.kv_len_max;
    sc.q_len_max = ri.num_tokens_in_batch > sc.q_len_max ? ri.num_tokens_in_batch : sc.q_len_max;
    sc.add(i, ri, ri.num_tokens_in_batch, tl, bc->causalMask[i].non_tree_cache_size, ri.prompt_phase);
  }
  h->num_generation_tokens_cur = bc->num_generation_tokens;
  rc = h->profiling == 2 ? FFB_OK : launch_prep(h, FFB_TREE_VERIFY_MODE, meta, qkv, T, bc->num_generation_tokens, 0, s);
  if (rc != FFB_OK) {
    return rc;
  }
  return launch_multi_dispatch(h, meta, out, T, sc.kv_len_max, sc.q_len_max, sc.reqs, sc.nreq, s);
}

int ffb_spec_attn_forward(ffb_attn_t h, ffb_beam_search_batch_config const *bbc,
                          ffb_combined_batch_config_meta const *meta, int shard_id, void const *qkv, void *out,
                          ffb_stream_t stream) {
  (void)shard_id;
  ffb_batch_config const *bc = (ffb_batch_config const *)bbc;
  int rc = check_common(h, bc, meta, qkv, out, FFB_BEAM_SEARCH_MODE);
  if (rc != FFB_OK) {
    return rc;
  }
  int const T = bc->num_tokens;
  if (T == 0) {
    return FFB_OK;
  }
  cudaStream_t s = (cudaStream_t)stream;
  BatchScan sc;
  for (int i = 0; i < h->d.max_requests; i++) {
    if (bc->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &ri = bc->requestsInfo[i];
    ffb_bit_mask const &bm = bc->causalMask[i];
    bool const speculative = ri.first_token_offset_in_batch < bc->num_generation_tokens;
    int const tl = speculative ? bm.non_tree_cache_size + bm.tree_size + bm.prompt_size - 1
                               : ri.first_token_depth_in_request + ri.num_tokens_in_batch;
    int const ql = speculative ? bbc->beamRequestsInfo[i].sub_request_num : ri.num_tokens_in_batch;
    FFB_REQUIRE(tl <= h->cache_len, "request %d: %d tokens exceed the cache length %d", i, tl, h->cache_len);
    sc.kv_len_max = tl > sc.kv_len_max ? tl : sc.kv_len_max;
    sc.q_len_max = ql > sc.q_len_max ? ql : sc.q_len_max;
    sc.add(i, ri, ql, tl, bm.non_tree_cache_size, !speculative);
  }
  h->num_generation_tokens_cur = bc->num_generation_tokens;
  rc = h->profiling == 2 ? FFB_OK : launch_prep(h, FFB_BEAM_SEARCH_MODE, meta, qkv, T, bc->num_generation_tokens, bbc->speculative_request_num, s);
  if (rc != FFB_OK) {
    return rc;
  }
  return launch_multi_dispatch(h, meta, out, T, sc.kv_len_max, sc.q_len_max, sc.reqs, sc.nreq, s);
}

int ffb_attn_read_kv(ffb_attn_t h, int request, int pos0, int n, void *k, void *v, ffb_stream_t stream) {
  FFB_REQUIRE(h && request >= 0 && request < h->d.max_requests && pos0 >= 0 && n >= 0 && pos0 + n <= h->cache_len,
              "kv range out of bounds");
  return launch_kv_io(h, request, pos0, n, k, v, false, (cudaStream_t)stream);
}
int ffb_attn_write_kv(ffb_attn_t h, int request, int pos0, int n, void const *k, void const *v, ffb_stream_t stream) {
  FFB_REQUIRE(h && request >= 0 && request < h->d.max_requests && pos0 >= 0 && n >= 0 && pos0 + n <= h->cache_len,
              "kv range out of bounds");
  return launch_kv_io(h, request, pos0, n, (void *)k, (void *)v, true, (cudaStream_t)stream);
}

} // extern "C"
