// host_tree.cu — host-side SpecInfer tree logic behind the C ABI (integer only, no CUDA call): the bit masks the
// tree-verify / beam attention kernels read, and the accept / commit walk over a verified tree.
// Replaces RequestManager::{initBitMask, updateBitMask, appendPendingRequest, appendBitMask, traverse_verify_tree}
// (src/runtime/request_manager.cc:2382-2741).  The reference sets mask bits one at a time in nested loops; a tree
// layer's bits are contiguous, so every update here is one OR of a range word.
#include "common.cuh"

namespace {

// bits [lo, hi) of a 64-bit word, 0 <= lo <= hi <= 64
inline unsigned long long bit_range(int lo, int hi) {
  unsigned long long const upto_hi = hi >= 64 ? ~0ULL : ((1ULL << hi) - 1ULL);
  unsigned long long const upto_lo = lo >= 64 ? ~0ULL : ((1ULL << lo) - 1ULL);
  return upto_hi & ~upto_lo;
}

} // namespace

extern "C" {

void ffb_bitmask_init(ffb_bit_mask *bm, int init_length) {
  // a fresh request: the whole prompt is one "layer" (request_manager.cc:2382-2396; the mask words are left alone)
  bm->non_tree_cache_size = 0;
  bm->tree_size = 1;
  bm->prompt_size = init_length;
  bm->this_layer_size = init_length;
}

void ffb_bitmask_update(ffb_bit_mask *bm, int init_length, int non_tree_size) {
  // after a verify step: everything accepted so far moves into the non-tree cache, the tree restarts at its root
  // (request_manager.cc:2399-2427: prompt_size = 1, so the causal triangle is the single bit mask[0] |= 1)
  bm->non_tree_cache_size = non_tree_size + init_length - 1;
  bm->tree_size = 1;
  bm->this_layer_size = init_length;
  bm->prompt_size = 1;
  bm->mask[0] |= 1ULL;
}

void ffb_bitmask_append_pending(ffb_bit_mask *bm, int init_length) {
  // another prompt chunk of a pending request (request_manager.cc:2430-2446)
  bm->non_tree_cache_size = 0;
  bm->tree_size = 1;
  bm->prompt_size += init_length;
  bm->this_layer_size = init_length;
}

int ffb_bitmask_append(ffb_bit_mask *bm, int new_nodes, int const *layer_nodes, int current_depth) {
  FFB_REQUIRE(bm && new_nodes >= 0 && current_depth >= 0 && (layer_nodes || current_depth <= 1), "bad argument");
  int const first_new = bm->tree_size, end_new = first_new + new_nodes;
  bm->tree_size = end_new;
  bm->this_layer_size = new_nodes;
  FFB_REQUIRE(end_new <= FFB_MAX_SPEC_TREE_TOKEN_NUM, "tree of %d nodes exceeds MAX_SPEC_TREE_TOKEN_NUM", end_new);
  FFB_REQUIRE(bm->prompt_size >= 0 && bm->prompt_size <= FFB_MAX_SPEC_TREE_TOKEN_NUM, "prompt_size %d out of range",
              bm->prompt_size);
  // every prompt position sees the whole new layer
  unsigned long long const layer_bits = bit_range(first_new, end_new);
  for (int i = 0; i < bm->prompt_size; i++) {
    bm->mask[i] |= layer_bits;
  }
  // ancestors: in every earlier tree layer the new nodes are dealt out in equal consecutive groups, one group per
  // node of that layer (request_manager.cc:2484-2503; the remainder of an uneven split belongs to nobody)
  int node = bm->prompt_size;
  for (int layer = 1; layer < current_depth; layer++) {
    int const width = layer_nodes[layer];
    if (width <= 0) {
      continue;
    }
    int const group = new_nodes / width;
    for (int j = 0; j < width; j++, node++) {
      if (node >= 0 && node < FFB_MAX_SPEC_TREE_TOKEN_NUM) {
        bm->mask[node] |= bit_range(first_new + j * group, first_new + (j + 1) * group);
      }
    }
  }
  if (node != first_new) {
    ffb::set_error("layer widths cover %d tree nodes, the mask holds %d", node, first_new);
    return FFB_ERR_STATE;
  }
  for (int i = first_new; i < end_new; i++) { // a node sees itself
    bm->mask[i] |= 1ULL << i;
  }
  return FFB_OK;
}

int ffb_traverse_verify_tree(ffb_tree_node const *input_tree, int n_in, ffb_tree_node const *llm_outputs, int n_out,
                             ffb_commit_entry const *committed_in, ffb_tree_node *verified, int *n_verified,
                             ffb_commit_entry *committed_out, int *n_committed) {
  FFB_REQUIRE(input_tree && llm_outputs && committed_in && verified && n_verified && committed_out && n_committed,
              "null argument");
  FFB_REQUIRE(n_in >= n_out && n_out >= 0, "input tree (%d nodes) smaller than the LLM output (%d)", n_in, n_out);
  int nv = 0, nc = 0;
  int layer_start = 0;    // index of the first node of the layer (run of equal depths) node i belongs to
  int accepted_slot = -1; // slot inside its layer of the first accepted non-root node; never reset (reference)
  for (int i = 0; i < n_out; i++) {
    if (i > 0 && input_tree[i].depth != input_tree[i - 1].depth) {
      layer_start = i;
    }
    bool take;
    if (i == 0) {
      take = true;
    } else {
      ffb_tree_node const &last = verified[nv - 1];
      if (input_tree[i].token_id != last.token_id || input_tree[i].depth != last.depth) {
        continue;
      }
      int const slot = i - layer_start;
      take = accepted_slot < 0 || slot == accepted_slot;
      if (accepted_slot < 0) {
        accepted_slot = slot;
      }
    }
    if (take) {
      verified[nv++] = llm_outputs[i];
      committed_out[nc].depth = input_tree[i].depth;
      committed_out[nc++].index_in_batch = committed_in[i].index_in_batch;
    }
    if (committed_in[i].depth != input_tree[i].depth) {
      ffb::set_error("committed depth %d of node %d disagrees with the tree depth %d", committed_in[i].depth, i,
                     input_tree[i].depth);
      return FFB_ERR_STATE;
    }
  }
  *n_verified = nv;
  *n_committed = nc;
  return FFB_OK;
}

} // extern "C"
