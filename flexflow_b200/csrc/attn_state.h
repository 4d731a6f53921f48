// attn_state.h — the state behind ffb_attn_t (role of IncMultiHeadSelfAttentionMeta,
// include/flexflow/ops/inc_multihead_self_attention.h:139-202).
#pragma once
#include "common.cuh"

struct ffb_attn_s {
  ffb_attn_desc d;
  int G;         // q heads per kv head
  int cache_len; // max_sequence_length (+ max_spec_tree_tokens outside incremental mode)
  int elt;       // bytes per element
  // Device memory, one allocation (the reference reserves one Realm instance, inc_…cu:1719-1760).
  // KV cache layout (private to the op): [request][kv_head][cache_len][head_dim] so a
  // head's rows are contiguous 2*head_dim-byte lines and only DISTINCT kv heads are stored.
  void *base = nullptr;
  size_t bytes = 0;
  void *k_cache = nullptr, *v_cache = nullptr;
  void *q_buf = nullptr;   // rotated queries [max_tokens][num_q_heads][head_dim]
  void *prev_kv = nullptr; // K,V of the previous batch [max_tokens][2][num_kv_heads][head_dim] (commit source)
  ffb::QueryDesc *qdesc = nullptr; // [max_tokens]
  int *split_counter = nullptr;    // [max_tokens*num_q_heads] arrival counters (self-resetting) for the fused combine
  float *partial = nullptr;        // split-KV partials
  size_t partial_bytes = 0;
  // m->profiling (inc_…cu:1510-1538): events around the dominant attention kernel only
  // 2 = measurement mode: the forward call launches ONLY that kernel, on the state the last full call left
  int profiling = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  int num_generation_tokens_cur = 0; // bc->num_generation_tokens of the batch in flight
  int num_active_infr_tokens = 0; // TreeIncMultiHeadSelfAttentionMeta::num_active_infr_tokens
  // TMA descriptors (CUtensorMap, 128 bytes each) over the K and V caches viewed as [rows][head_dim] (attn_tc.cu)
  alignas(64) unsigned char tmK[128];
  alignas(64) unsigned char tmV[128];
  bool tm_ready = false;
};

namespace ffb {

// launch wrappers implemented in the .cu files ------------------------------------------------
// prep: RoPE + query staging + KV append + per-query descriptors, one pass over [T][3][H]
int launch_prep(ffb_attn_s *h, int mode, ffb_combined_batch_config_meta const *meta, void const *qkv,
                int T, int num_generation_tokens, int speculative_request_num, cudaStream_t s);
// commit: cache[req][token_depth] <- prev_kv[token_index]  (tree_…cu:335-396)
int launch_commit(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, int n_commit,
                  cudaStream_t s);
// decode attention (one query per (token, head)), split-KV + combine
int launch_decode_attn(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T,
                       int kv_len_max, cudaStream_t s);
// multi-query (tree / beam / prompt / GQA) attention
int launch_multi_attn(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T,
                      int kv_len_max, int q_len_max, cudaStream_t s);
// tcgen05 / TMA version of the multi-query kernel (attn_tc.cu)
int attn_tc_init(ffb_attn_s *h);
int attn_tc_supported(ffb_attn_s *h);
// one active request as the host sees it (from the BatchConfig), input of the tcgen05 kernel's stream-K schedule
struct AtcReq {
  int slot, first_tok, qlen; // request slot, first_token_offset_in_batch, query tokens in this batch
  int tl;                    // keys visible to the request's last query token
  int ntcs;                  // causalMask[slot].non_tree_cache_size
  int causal, first_depth;   // causal != 0: token qi sees keys [0, first_depth + qi + 1)
};
int launch_attn_tc(ffb_attn_s *h, ffb_combined_batch_config_meta const *meta, void *out, int T, AtcReq const *reqs,
                   int nreq, cudaStream_t s);
int launch_kv_io(ffb_attn_s *h, int request, int pos0, int n, void *k, void *v, bool write,
                 cudaStream_t s);

} // namespace ffb
