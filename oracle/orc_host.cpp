// orc_host.cpp — restatement of the host-side SpecInfer tree logic whose integer outputs feed the
// attention kernels (bit masks) and come out of verification (accepted tokens, commit list).
// TEST INFRASTRUCTURE ONLY (see orc_common.h).  Integer-only: results must be bit-exact.
#include "orc_common.h"

extern "C" {

// RequestManager::initBitMask — src/runtime/request_manager.cc:2382-2396
void orc_init_bitmask(ffb_bit_mask *bm, int initLength) {
  bm->non_tree_cache_size = 0;
  bm->tree_size = 1;
  bm->prompt_size = initLength;
  bm->this_layer_size = initLength;
}

// RequestManager::updateBitMask — request_manager.cc:2399-2427.  `wide` = 1 uses the intended
// 64-bit shift (1ULL << j); wide = 0 reproduces the reference's 32-bit `1 << j`,
// identical for j <= 30.
void orc_update_bitmask(ffb_bit_mask *bm, int initLength, int non_tree_size, int wide) {
  bm->non_tree_cache_size = non_tree_size + initLength - 1;
  bm->tree_size = 1;
  bm->this_layer_size = initLength;
  bm->prompt_size = 1;
  for (int i = 0; i < bm->prompt_size; i++) {
    for (int j = i; j < bm->prompt_size; j++) {
      bm->mask[i] |= wide ? (1ULL << j) : (unsigned long long)(long long)(1 << j);
    }
  }
}

// RequestManager::appendPendingRequest — request_manager.cc:2430-2446
void orc_append_pending_request(ffb_bit_mask *bm, int initLength) {
  bm->non_tree_cache_size = 0;
  bm->tree_size = 1;
  bm->prompt_size += initLength;
  bm->this_layer_size = initLength;
}

// RequestManager::appendBitMask — request_manager.cc:2448-2517.  layer_nodes[i] restates
// tree.treeLayers[i].nodes_num_this_layer for i in [0, currentDepth).
int orc_append_bitmask(ffb_bit_mask *bm, int newNodes, int const *layer_nodes, int currentDepth,
                       int wide) {
  auto bit = [&](int j) { return wide ? (1ULL << j) : (unsigned long long)(long long)(1 << j); };
  int pre_tree_size = bm->tree_size;
  bm->tree_size += newNodes;
  bm->this_layer_size = newNodes;
  if (bm->tree_size > FFB_MAX_SPEC_TREE_TOKEN_NUM) {
    return 1;
  }
  for (int i = 0; i < bm->prompt_size; i++) {
    for (int j = pre_tree_size; j < bm->tree_size; j++) {
      bm->mask[i] |= bit(j);
    }
  }
  int token_idx = bm->prompt_size;
  int new_nodes_start_idx = pre_tree_size;
  for (int i = 1; i < currentDepth; i++) {
    new_nodes_start_idx = pre_tree_size;
    int nodes_this_layer = layer_nodes[i];
    for (int j = 0; j < nodes_this_layer; j++) {
      int group_size = newNodes / nodes_this_layer;
      for (int k = 0; k < group_size; k++) {
        bm->mask[token_idx] |= bit(new_nodes_start_idx);
        new_nodes_start_idx += 1;
      }
      token_idx += 1;
    }
  }
  if (token_idx != pre_tree_size) {
    return 2; // assert(token_idx == pre_tree_size)
  }
  for (int i = token_idx; i < bm->tree_size; i++) {
    bm->mask[i] |= bit(i);
  }
  return 0;
}

// RequestManager::traverse_verify_tree — request_manager.cc:2583-2741.
// Trees are serialised in layer order as (token id, depth) pairs.  `committed_*` is the request's
// committed_tokens list on entry, (depth, index_in_batch) pairs.  Outputs: verified (token, depth)
// pairs and the new committed (depth, index_in_batch) list.  Returns the number of verified tokens,
// n_committed through *n_committed_out; negative on a violated reference assert.
int orc_traverse_verify_tree(int n_in, int const *in_tok, int const *in_depth, int n_out,
                             int const *out_tok, int const *out_depth, int const *committed_depth,
                             int const *committed_index, int *ver_tok, int *ver_depth,
                             int *new_committed_depth, int *new_committed_index,
                             int *n_committed_out) {
  if (n_in < n_out) {
    return -1;
  }
  std::vector<int> treeLayers(n_in > 0 ? n_in : 1);
  int node_num = 1, layer_num = 0;
  for (int t = 0; t < n_in; t++) {
    if (t == n_in - 1 || in_depth[t + 1] != in_depth[t]) {
      treeLayers[layer_num] = node_num;
      layer_num += 1;
      node_num = 1;
    } else {
      node_num++;
    }
  }
  bool findFirst = false;
  layer_num = -1;
  int first_layer_slot = 0;
  int processed_whole_layer_tokens = 0;
  int nv = 0, nc = 0;
  for (int i = 0; i < n_out; i++) {
    if (i == 0 || in_depth[i - 1] != in_depth[i]) {
      layer_num += 1;
      processed_whole_layer_tokens += i == 0 ? 0 : treeLayers[layer_num - 1];
    }
    if (i == 0) {
      ver_tok[nv] = out_tok[i];
      ver_depth[nv++] = out_depth[i];
      new_committed_depth[nc] = in_depth[i];
      new_committed_index[nc++] = committed_index[i];
      if (committed_depth[i] != in_depth[i]) {
        return -2;
      }
      continue;
    }
    if (in_tok[i] == ver_tok[nv - 1] && in_depth[i] == ver_depth[nv - 1]) {
      if (findFirst) {
        int layer_slot = i - processed_whole_layer_tokens;
        if (first_layer_slot == layer_slot) {
          ver_tok[nv] = out_tok[i];
          ver_depth[nv++] = out_depth[i];
          new_committed_depth[nc] = in_depth[i];
          new_committed_index[nc++] = committed_index[i];
        }
      } else {
        ver_tok[nv] = out_tok[i];
        ver_depth[nv++] = out_depth[i];
        first_layer_slot = i - processed_whole_layer_tokens;
        findFirst = true;
        new_committed_depth[nc] = in_depth[i];
        new_committed_index[nc++] = committed_index[i];
      }
      if (committed_depth[i] != in_depth[i]) {
        return -2;
      }
    }
  }
  *n_committed_out = nc;
  return nv;
}

// Kernels::Linear forward — src/ops/kernels/linear_kernels.cu:450-582: out[b][o] = sum_i in[b][i]*W[o][i]
// (+ bias).  fp32 accumulation of DT inputs, result stored as DT (the reference accumulates in DT
// inside cuBLAS; tolerance covers it).
void orc_linear_forward(float const *in, float const *w, float const *bias, float *out, int batch,
                        int in_dim, int out_dim, int dt) {
  orc_parallel_for(out_dim, [&](int o) {
    for (int b = 0; b < batch; b++) {
      double acc = 0.0;
      float const *x = in + (size_t)b * in_dim, *wr = w + (size_t)o * in_dim;
      for (int i = 0; i < in_dim; i++) {
        acc += (double)x[i] * (double)wr[i];
      }
      float v = (float)acc;
      if (bias) {
        v = round_dt(round_dt(v, dt) + bias[o], dt);
      }
      out[(size_t)b * out_dim + o] = round_dt(v, dt);
    }
  });
}

} // extern "C"
