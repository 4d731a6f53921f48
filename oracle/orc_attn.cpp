// orc_attn.cpp — CPU restatement of the three attention operators' device code.
// TEST INFRASTRUCTURE ONLY (see orc_common.h).  Layouts follow the reference exactly:
//   qkv   [T][3][H]            H = num_q_heads*head_dim, K/V at q-head granularity
//   cache [max_requests][max_len][H]   (src/ops/inc_multihead_self_attention.cu:1693-1712)
//   out   [T][H]
#include "orc_common.h"

namespace {

struct OrcAttn {
  int mode, dt, nq, dh, H;
  int max_requests, max_seq, max_spec, max_tokens, cache_len;
  int apply_rope, llama3, orig_max_pos;
  float theta, factor, low_ff, high_ff;
  int scaling_query, qk_prod_scaling, position_bias, global_nq;
  float scaling_factor;
  // bit 0: keep the reference's running qk_max across query tokens (tree/spec kernels declare it
  //        outside the qi loop); bit 1: keep the spec prompt loop's skip of prompt requests while
  //        tokens_previous_requests < num_generation_tokens (spec_…cu:489-492), which leaves those
  //        tokens' outputs uncomputed in a mixed beam batch.
  int ref_quirks;
  int num_active_infr_tokens = 0; // TreeIncMultiHeadSelfAttentionMeta::num_active_infr_tokens
  std::vector<float> qkv;         // devQKVProjArray [max_tokens][3][H] (persists across calls)
  std::vector<float> kcache, vcache;
  std::vector<float> attn_heads;
  size_t cache_index(int req, int pos, int off) const {
    return ((size_t)req * cache_len + pos) * H + off;
  }
};

// apply_rotary_embedding_hf — src/ops/inc_multihead_self_attention.cu:664-738.
// One "thread" i per complex pair of Q (first T*H/2) and K (second T*H/2).
void rope_hf(OrcAttn &m, ffb_per_token_info const *tokenInfos, int T) {
  int H = m.H, proj = m.dh;
  size_t q_array_size = (size_t)proj * T * m.nq;
  for (size_t i = 0; i < (size_t)T * H; i++) {
    bool q_tensor = i < q_array_size / 2;
    size_t real_i = q_tensor ? i : i - q_array_size / 2;
    int token_idx = (int)(real_i / (H / 2));
    int idx = (int)(real_i % (proj / 2));
    int head_idx = (int)((real_i - (size_t)token_idx * (H / 2)) / (proj / 2));
    size_t re = idx + (size_t)head_idx * proj + (size_t)token_idx * H * 3 + (size_t)H * (q_tensor ? 0 : 1);
    size_t im = re + proj / 2;
    float x = m.qkv[re], y = m.qkv[im];
    size_t pos = (size_t)tokenInfos[token_idx].abs_depth_in_request;
    int pos_i = (int)(real_i % (proj / 2));
    // float freq = pos * (1.0 / pow(rope_theta, (float)2 * pos_i / proj_size));   (:708-709)
    float freq = (float)((double)pos * (1.0 / (double)powf(m.theta, (float)2 * pos_i / proj)));
    if (m.llama3) { // :711-730
      float pi = 3.14159265358979323846f;
      float wavelen = 2 * pi / freq;
      float low_freq_wavelen = m.orig_max_pos / m.low_ff;
      float high_freq_wavelen = m.orig_max_pos / m.high_ff;
      if (wavelen < high_freq_wavelen) {
      } else if (wavelen > low_freq_wavelen) {
        freq = freq / m.factor;
      } else {
        float smooth = (m.orig_max_pos / wavelen - m.low_ff) / (m.high_ff - m.low_ff);
        freq = ((1 - smooth) * freq / m.factor + smooth * freq);
      }
    }
    float c = cosf(freq), s = sinf(freq);
    // cuCmulf: (x + iy)(c + is)
    m.qkv[re] = round_dt(x * c - y * s, m.dt);
    m.qkv[im] = round_dt(x * s + y * c, m.dt);
  }
}

// compute_qkv_kernel — inc_multihead_self_attention.cu:807-857 (scaling_query_kernel :650-662)
void compute_qkv(OrcAttn &m, ffb_per_token_info const *tokenInfos, int T) {
  if (m.scaling_query) {
    for (int t = 0; t < T; t++) {
      for (int o = 0; o < m.H; o++) {
        float &v = m.qkv[(size_t)t * 3 * m.H + o];
        v = round_dt(v * m.scaling_factor, m.dt);
      }
    }
  }
  if (m.apply_rope) {
    rope_hf(m, tokenInfos, T);
  }
}

// One softmax(QK^T)V row, exactly in the order of the reference generation kernels
// (inc_…cu:372-623): logits = scale*dot (masked -> 0 and excluded from the max), p = exp(l - max),
// inv = 1/(sum + 1e-6), out = sum p*inv*v accumulated in fp32, stored as DT.
// `visible(ti)` is the complement of the reference's `mask`.  qk_max is in/out so callers can
// reproduce the running maximum of the tree/spec kernels.
template <typename Vis>
void attend_row(OrcAttn const &m, float const *q, int req, int head, int tlength, float scale,
                Vis visible, float &qk_max, float *out) {
  std::vector<float> logits(tlength);
  for (int ti = 0; ti < tlength; ti++) {
    float const *k = &m.kcache[m.cache_index(req, ti, head * m.dh)];
    float dot = 0.f;
    for (int d = 0; d < m.dh; d++) {
      dot += q[d] * k[d];
    }
    float qk = scale * dot;
    bool mask = !visible(ti);
    qk_max = mask ? qk_max : fmaxf(qk_max, qk);
    logits[ti] = mask ? 0.f : qk;
  }
  float exp_sum = 0.f;
  for (int ti = 0; ti < tlength; ti++) {
    float logit = visible(ti) ? expf(logits[ti] - qk_max) : 0.f;
    exp_sum += logit;
    logits[ti] = logit;
  }
  float inv_sum = 1.f / (exp_sum + 1.e-6f);
  std::vector<float> acc(m.dh, 0.f);
  for (int ti = 0; ti < tlength; ti++) {
    if (!visible(ti)) {
      continue;
    }
    float p = logits[ti] * inv_sum;
    float const *v = &m.vcache[m.cache_index(req, ti, head * m.dh)];
    for (int d = 0; d < m.dh; d++) {
      acc[d] += p * v[d];
    }
  }
  for (int d = 0; d < m.dh; d++) {
    out[d] = round_dt(acc[d], m.dt);
  }
}

float attn_scale(OrcAttn const &m) { return m.qk_prod_scaling ? 1.0f / sqrtf((float)m.dh) : 1.0f; }

} // namespace

namespace {

// compute_attention_kernel_prompt — inc_multihead_self_attention.cu:98-366 (spec twin
// spec_inc_multihead_self_attention.cu:461-679).  cuBLAS/cuDNN are closed libraries, so this is
// the published math of the three calls: S = alpha*Q K^T stored as DT, ALiBi bias (:213-226,
// apply_position_bias_qkprd :625-648), causal -inf fill, softmax over keys (CUDNN_SOFTMAX_ACCURATE
// = max-subtracted), O = P V stored as DT.  Causal rule is the INTENDED one
// (key <= total-new+query, spec_fill_entries_above_diagonal spec_…cu:444-459); the inc variant's
// fill kernel (:79-96) indexes the matrix with the wrong leading dimension when total > new
// (chunked prefill) — see DESIGN.md "reference quirks".
void prompt_request(OrcAttn &m, int req, int first_offset, int num_new, int total, int shard_id) {
  float alpha = round_dt(attn_scale(m), m.dt);
  std::vector<float> s(total);
  for (int h = 0; h < m.nq; h++) {
    for (int qi = 0; qi < num_new; qi++) {
      float const *q = &m.qkv[(size_t)(first_offset + qi) * 3 * m.H + h * m.dh];
      float mx = -INFINITY;
      int limit = total - num_new + qi;
      for (int k = 0; k < total; k++) {
        float const *kk = &m.kcache[m.cache_index(req, k, h * m.dh)];
        float dot = 0.f;
        for (int d = 0; d < m.dh; d++) {
          dot += q[d] * kk[d];
        }
        float v = round_dt(alpha * dot, m.dt);
        if (m.position_bias) {
          int head_idx = h + m.nq * shard_id;
          float base = (float)(head_idx + 1) * 8 / m.global_nq;
          float slopes = (float)(1.0 / pow(2, base));
          v = round_dt(v + round_dt((k + 1 - total) * slopes, m.dt), m.dt);
        }
        if (k > limit) {
          v = -INFINITY;
        }
        s[k] = v;
        mx = std::max(mx, v);
      }
      float sum = 0.f;
      for (int k = 0; k < total; k++) {
        s[k] = expf(s[k] - mx);
        sum += s[k];
      }
      std::vector<float> acc(m.dh, 0.f);
      for (int k = 0; k < total; k++) {
        float p = round_dt(s[k] / sum, m.dt);
        float const *vv = &m.vcache[m.cache_index(req, k, h * m.dh)];
        for (int d = 0; d < m.dh; d++) {
          acc[d] += p * vv[d];
        }
      }
      float *o = &m.attn_heads[(size_t)(first_offset + qi) * m.H + h * m.dh];
      for (int d = 0; d < m.dh; d++) {
        o[d] = round_dt(acc[d], m.dt);
      }
    }
  }
}

inline bool bit_visible(ffb_bit_mask const &bm, int ti, int bit, bool quirk32) {
  if (ti < bm.non_tree_cache_size) {
    return true;
  }
  // Reference: mask[...] & (1 << qi) with a 32-bit `1` — identical to 1ULL for bit <= 30.
  // The build uses the intended 64-bit semantics; quirk32 reproduces the reference for bit <= 30.
  (void)quirk32;
  return (bm.mask[ti - bm.non_tree_cache_size] >> bit) & 1ULL;
}

void load_qkv(OrcAttn &m, float const *qkv, int T) {
  std::memcpy(m.qkv.data(), qkv, sizeof(float) * (size_t)T * 3 * m.H); // phase 0 D2D copy
}

} // namespace

extern "C" {

void *orc_attn_create(int mode, int dt, int nq, int dh, int max_requests, int max_seq, int max_spec,
                      int max_tokens, int apply_rope, float theta, int llama3, float factor,
                      float low_ff, float high_ff, int orig_max_pos, int scaling_query,
                      float scaling_factor, int qk_prod_scaling, int position_bias, int global_nq) {
  OrcAttn *m = new OrcAttn();
  m->mode = mode; m->dt = dt; m->nq = nq; m->dh = dh; m->H = nq * dh;
  m->max_requests = max_requests; m->max_seq = max_seq; m->max_spec = max_spec;
  m->max_tokens = max_tokens;
  // inc_…cu:1691-1716: tree/beam modes add max_spec_tree_token_num rows
  m->cache_len = max_seq + (mode == FFB_INC_DECODING_MODE ? 0 : max_spec);
  m->apply_rope = apply_rope; m->theta = theta; m->llama3 = llama3; m->factor = factor;
  m->low_ff = low_ff; m->high_ff = high_ff; m->orig_max_pos = orig_max_pos;
  m->scaling_query = scaling_query; m->scaling_factor = scaling_factor;
  m->qk_prod_scaling = qk_prod_scaling; m->position_bias = position_bias; m->global_nq = global_nq;
  m->ref_quirks = 3;
  m->qkv.assign((size_t)max_tokens * 3 * m->H, 0.f);
  m->attn_heads.assign((size_t)max_tokens * m->H, 0.f);
  m->kcache.assign((size_t)max_requests * m->cache_len * m->H, 0.f);
  m->vcache.assign((size_t)max_requests * m->cache_len * m->H, 0.f);
  return m;
}
void orc_attn_destroy(void *p) { delete (OrcAttn *)p; }
float *orc_attn_kcache(void *p) { return ((OrcAttn *)p)->kcache.data(); }
float *orc_attn_vcache(void *p) { return ((OrcAttn *)p)->vcache.data(); }
int orc_attn_cache_len(void *p) { return ((OrcAttn *)p)->cache_len; }
void orc_attn_set_quirks(void *p, int q) { ((OrcAttn *)p)->ref_quirks = q; }

static int num_active_tokens(ffb_batch_config const *bc) { return bc->num_tokens; }

// Kernels::IncMultiHeadAttention::inference_kernel — inc_multihead_self_attention.cu:944-985
int orc_inc_attn_forward(void *p, ffb_batch_config const *bc, int shard_id, float const *qkv,
                         float *out) {
  OrcAttn &m = *(OrcAttn *)p;
  int T = num_active_tokens(bc);
  load_qkv(m, qkv, T);
  compute_qkv(m, bc->tokensInfo, T);
  // store_kv_cache :35-61
  for (int t = 0; t < T; t++) {
    int req = bc->tokensInfo[t].request_index, pos = bc->tokensInfo[t].abs_depth_in_request;
    for (int o = 0; o < m.H; o++) {
      m.kcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + m.H + o];
      m.vcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + 2 * m.H + o];
    }
  }
  float scale = attn_scale(m);
  // compute_attention_kernel_generation :902-927, kernel :372-623: grid (heads, generation tokens)
  orc_parallel_for(bc->num_generation_tokens * m.nq, [&](int w) {
    int request_idx = w / m.nq, h = w % m.nq;
    int slot = bc->requestsInfo[request_idx].batch_config_request_id;
    int tlength = bc->requestsInfo[slot].first_token_depth_in_request +
                  bc->requestsInfo[slot].num_tokens_in_batch;
    float qk_max = -FLT_MAX;
    attend_row(m, &m.qkv[(size_t)request_idx * 3 * m.H + h * m.dh], slot, h, tlength, scale,
               [](int) { return true; }, qk_max, &m.attn_heads[(size_t)request_idx * m.H + h * m.dh]);
  });
  if (bc->num_tokens > bc->num_generation_tokens) { // :974-977
    for (int i = 0; i < m.max_requests; i++) {
      if (bc->request_completed[i] || !bc->requestsInfo[i].prompt_phase) {
        continue;
      }
      int num_new = bc->requestsInfo[i].num_tokens_in_batch;
      int total = bc->requestsInfo[i].first_token_depth_in_request + num_new;
      prompt_request(m, i, bc->requestsInfo[i].first_token_offset_in_batch, num_new, total, shard_id);
    }
  }
  std::memcpy(out, m.attn_heads.data(), sizeof(float) * (size_t)T * m.H);
  return 0;
}

} // extern "C"

extern "C" {

// Kernels::TreeIncMultiHeadAttention::inference_kernel — tree_inc_multihead_self_attention.cu:575-626
int orc_tree_attn_forward(void *p, ffb_tree_verify_batch_config const *bc, int shard_id,
                          float const *qkv, float *out) {
  OrcAttn &m = *(OrcAttn *)p;
  (void)shard_id;
  // commit_tokens_kernel :335-370 — reads the PREVIOUS batch's devQKVProjArray
  for (int c = 0; c < bc->num_tokens_to_commit; c++) {
    int src = bc->committed_tokens[c].token_index;
    if (src >= m.num_active_infr_tokens) {
      return 5; // assert(token_idx_in_last_batch < num_active_tokens_in_last_batch) :353
    }
    int req = bc->committed_tokens[c].request_index, pos = bc->committed_tokens[c].token_depth;
    for (int o = 0; o < m.H; o++) {
      m.kcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)src * 3 * m.H + m.H + o];
      m.vcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)src * 3 * m.H + 2 * m.H + o];
    }
  }
  int T = bc->num_tokens;
  m.num_active_infr_tokens = T; // :594
  load_qkv(m, qkv, T);
  compute_qkv(m, bc->tokensInfo, T);
  // update_tree_branch_kv_cache_fused :433-478 — slot = position in batch, not abs_depth
  for (int t = 0; t < T; t++) {
    int req = bc->tokensInfo[t].request_index;
    int pos = t + bc->requestsInfo[req].first_token_depth_in_request -
              bc->requestsInfo[req].first_token_offset_in_batch;
    for (int o = 0; o < m.H; o++) {
      m.kcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + m.H + o];
      m.vcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + 2 * m.H + o];
    }
  }
  float scale = attn_scale(m);
  // compute_attention_kernel_fused_kernel :35-333: grid (heads, num_active_requests)
  int R = 0;
  for (int i = 0; i < m.max_requests; i++) {
    R += bc->request_completed[i] ? 0 : 1;
  }
  for (int request_idx = 0; request_idx < R; request_idx++) {
    int slot = bc->requestsInfo[request_idx].batch_config_request_id;
    ffb_per_request_info const &ri = bc->requestsInfo[slot];
    int tlength = ri.first_token_depth_in_request + ri.num_tokens_in_batch;
    int qlength = ri.num_tokens_in_batch;
    ffb_bit_mask const &bm = bc->causalMask[slot];
    int first_token_idx = 0; // :92-96
    for (int r = 0; r < slot; r++) {
      first_token_idx += bc->request_completed[r] ? 0 : bc->requestsInfo[r].num_tokens_in_batch;
    }
    bool prompt_phase = ri.prompt_phase;
    int q_start = ri.first_token_depth_in_request;
    for (int h = 0; h < m.nq; h++) {
      float qk_max = -FLT_MAX; // declared OUTSIDE the qi loop in the reference (:107)
      for (int qi = 0; qi < qlength; qi++) {
        if (!(m.ref_quirks & 1)) {
          qk_max = -FLT_MAX;
        }
        auto vis = [&](int ti) {
          return prompt_phase ? !(qi + q_start < ti) : bit_visible(bm, ti, qi, true);
        };
        attend_row(m, &m.qkv[(size_t)(first_token_idx + qi) * 3 * m.H + h * m.dh], slot, h, tlength,
                   scale, vis, qk_max, &m.attn_heads[(size_t)(first_token_idx + qi) * m.H + h * m.dh]);
      }
    }
  }
  std::memcpy(out, m.attn_heads.data(), sizeof(float) * (size_t)T * m.H);
  return 0;
}

// Kernels::SpecIncMultiHeadSelfAttention::inference_kernel — spec_inc_multihead_self_attention.cu:681-724
int orc_spec_attn_forward(void *p, ffb_beam_search_batch_config const *bc, int shard_id,
                          float const *qkv, float *out) {
  OrcAttn &m = *(OrcAttn *)p;
  int T = bc->num_tokens;
  load_qkv(m, qkv, T);
  compute_qkv(m, bc->tokensInfo, T);
  // spec_inc_store_kv_cache :311-358
  for (int t = 0; t < T; t++) {
    int req = bc->tokensInfo[t].request_index;
    ffb_bit_mask const &bm = bc->causalMask[req];
    int pos = bm.prompt_size + bm.non_tree_cache_size + bm.tree_size - 1 - bm.this_layer_size + t -
              bc->requestsInfo[req].first_token_offset_in_batch;
    for (int o = 0; o < m.H; o++) {
      m.kcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + m.H + o];
      m.vcache[m.cache_index(req, pos, o)] = m.qkv[(size_t)t * 3 * m.H + 2 * m.H + o];
    }
  }
  float scale = attn_scale(m);
  if (bc->num_generation_tokens > 0) {
    // compute_spec_inc_attention_kernel_generation_kernel :36-309: grid (heads, speculative_request_num)
    for (int request_idx = 0; request_idx < bc->speculative_request_num; request_idx++) {
      int slot = bc->requestsInfo[request_idx].batch_config_request_id;
      ffb_bit_mask const &bm = bc->causalMask[slot];
      int totalCacheSize = bm.non_tree_cache_size + bm.tree_size + bm.prompt_size - 1;
      int first_token_idx = 0; // :92-95
      for (int r = 0; r < slot; r++) {
        first_token_idx += bc->request_completed[r] ? 0 : bc->causalMask[r].this_layer_size;
      }
      int tree_branch_num = bc->beamRequestsInfo[slot].sub_request_num;
      for (int h = 0; h < m.nq; h++) {
        float qk_max = -FLT_MAX; // outside the qi loop (:106)
        for (int qi = 0; qi < tree_branch_num; qi++) {
          if (!(m.ref_quirks & 1)) {
            qk_max = -FLT_MAX;
          }
          int query_token = bm.prompt_size + bm.tree_size - 1 - tree_branch_num + qi; // :143-144
          auto vis = [&](int ti) { return bit_visible(bm, ti, query_token, true); };
          attend_row(m, &m.qkv[(size_t)(first_token_idx + qi) * 3 * m.H + h * m.dh], slot, h,
                     totalCacheSize, scale, vis, qk_max,
                     &m.attn_heads[(size_t)(first_token_idx + qi) * m.H + h * m.dh]);
        }
      }
    }
  }
  if (bc->num_tokens > bc->num_generation_tokens) { // spec_…cu:461-679
    int tokens_previous_requests = 0;
    for (int i = 0; i < m.max_requests; i++) {
      if (bc->request_completed[i] || !bc->requestsInfo[i].prompt_phase ||
          bc->requestsInfo[i].num_tokens_in_batch == 0) {
        continue;
      } else if ((m.ref_quirks & 2) && tokens_previous_requests < bc->num_generation_tokens) {
        tokens_previous_requests += bc->requestsInfo[i].num_tokens_in_batch;
        continue;
      }
      int num_new = bc->requestsInfo[i].num_tokens_in_batch;
      int total = bc->requestsInfo[i].first_token_depth_in_request + num_new;
      prompt_request(m, i, bc->requestsInfo[i].first_token_offset_in_batch, num_new, total, shard_id);
      tokens_previous_requests += num_new;
    }
  }
  std::memcpy(out, m.attn_heads.data(), sizeof(float) * (size_t)T * m.H);
  return 0;
}

} // extern "C"
