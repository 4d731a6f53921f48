// request_manager.cu — the batch builders of FlexFlow Serve's RequestManager behind the C ABI (host code, integer only):
// incremental decoding (prepare_next_batch, src/runtime/request_manager.cc:713-1138) and the three SpecInfer builders
// (prepare_next_batch_init :1170-1580, _beam :1610-1892 with store/update_beam_metadata :2217-2379, _verify :1923-2215
// with traverse_beam_tree / merge_dfs_trees :2743-2888).  They emit the reference's descriptors field for field
// (tests/test_request_manager.py runs them in lockstep with the reference's own functions, compiled from the reference
// tree by the test infrastructure), but are organised differently: requests live in a flat table indexed by guid - first
// guid, a request's speculation tree is a short array of layers, descriptors are zero-filled PODs, every reference
// assert is a returned status.  Also here: the D2H that ends a model step (InferenceResult / BeamInferenceResult).
#include <algorithm>
#include <cstring>
#include <deque>
#include <new>
#include <vector>

#include "common.cuh"

namespace {

constexpr size_t kFirstGuid = 1000000; // RequestManager::RequestManager(), request_manager.cc:102-106
constexpr int kLayers = FFB_MAX_BEAM_DEPTH + 1;
constexpr int kBranches = FFB_MAX_SPECULATIVE_TREE_BRANCHES;

struct TreeLayer { // BeamTree::treeLayer, include/flexflow/request_manager.h:103-111
  int count = 0;
  int token[kBranches] = {0, 0, 0};
  float prob[kBranches] = {0.f, 0.f, 0.f};
  int parent[kBranches] = {0, 0, 0};
};

struct Req {
  int status = FFB_REQ_PENDING;
  int max_length = -1;
  int initial_len = 0;
  int ssm_cache_size = 0, llm_cache_size = 0;
  int llm_steps = 0, ssm_steps = 0;
  std::vector<int> tokens;
  std::vector<TreeLayer> spec;           // [ssm][layer], beam_trees
  std::vector<ffb_tree_node> tree_in;    // dfs_tree_inputs[guid]: what the last verify batch carried for this request
  std::vector<ffb_commit_entry> commits; // committed_tokens[guid]
  bool has_commits = false;              // committed_tokens.count(guid) != 0
};

} // namespace

struct ffb_rm_s {
  ffb_rm_desc d;
  std::vector<int> widths; // spec_infer_tree_width
  std::vector<Req> reqs;
  std::deque<size_t> pending;
  size_t processed = 0;

  Req *find(size_t guid) {
    return guid >= kFirstGuid && guid - kFirstGuid < reqs.size() ? &reqs[guid - kFirstGuid] : nullptr;
  }
  int max_verify_tokens() const { return d.max_tokens_per_batch + d.max_spec_tree_token_num * d.max_requests_per_batch; }
  int width_at(int ssm_step) const { return (int)widths.size() > ssm_step ? widths[ssm_step] : 1; }
  bool is_eos(int tok) const {
    for (int i = 0; i < d.num_eos_tokens; i++) {
      if (d.eos_token_ids[i] == tok) {
        return true;
      }
    }
    return false;
  }
  bool finished_decoding(Req const &r, int max_length) const { // check_inf_req_completion, :645-657
    return (int)r.tokens.size() >= max_length || is_eos(r.tokens.back());
  }
};

namespace {

#define RM_STATE(cond, ...)              \
  do {                                   \
    if (!(cond)) {                       \
      ffb::set_error(__VA_ARGS__);       \
      return FFB_ERR_STATE;              \
    }                                    \
  } while (0)

// a default-constructed descriptor (BatchConfig::BatchConfig, src/runtime/batch_config.cc:57-70; PerRequestInfo's
// constructor, include/flexflow/batch_config.h:87-99): everything zero except request_completed = true,
// batch_config_request_id = -1, optimizer_tasks.compute_gradients = true
template <typename BC>
void fresh(BC *bc) {
  std::memset(bc, 0, sizeof(BC));
  for (int i = 0; i < FFB_MAX_NUM_REQUESTS; i++) {
    bc->request_completed[i] = true;
    bc->requestsInfo[i].batch_config_request_id = -1;
    bc->requestsInfo[i].optimizer_tasks[0] = true;
  }
}
void fresh_beam(ffb_beam_search_batch_config *bc) {
  fresh(bc);
  bc->beam_width = 1;        // DEFAULT_BEAM_WIDTH, src/runtime/beam_search_batch_config.cc:22
  bc->target_iterations = 3; // DEFAULT_TARGET_ITERATIONS
  for (int i = 0; i < FFB_MAX_NUM_REQUESTS; i++) {
    bc->beamRequestsInfo[i].current_depth = -1;
    bc->beamRequestsInfo[i].max_depth = FFB_MAX_BEAM_DEPTH;
  }
}

template <typename BC>
inline void put_token(BC *bc, int slot, int depth, int token) {
  ffb_per_token_info &t = bc->tokensInfo[bc->num_tokens++];
  t.request_index = slot;
  t.abs_depth_in_request = depth;
  t.token_id = token;
}

inline void root_beam_state(ffb_beam_per_request_info &b, int beam_size, int max_depth) {
  b.current_depth = 1;
  b.beam_size = beam_size;
  b.max_depth = max_depth;
  b.sub_request_num = 1;
  for (int j = 0; j < kBranches; j++) {
    b.parent_id[j] = 0;
    b.probs[j] = 1.f;
  }
}

// store_beam_metadata (request_manager.cc:2217-2325): file the SSM's candidates of this step as the next layer of each
// request's tree.  Results are laid out request after request in batch order; a request's candidates sit behind
// beam_size entries for each of its earlier positions.
int record_ssm_step(ffb_rm_s *rm, ffb_beam_search_batch_config const *bc, ffb_beam_inference_result const *res) {
  int const T = bc->num_tokens;
  if (T <= 0) {
    return FFB_OK;
  }
  int const ssm = bc->model_id;
  int cursor = 0;
  int first = 0; // first token of the current request's run
  for (int t = 1; t <= T; t++) {
    int const slot = bc->tokensInfo[first].request_index;
    if (t < T && bc->requestsInfo[bc->tokensInfo[t].request_index].request_guid == bc->requestsInfo[slot].request_guid) {
      continue;
    }
    int const last_slot = bc->tokensInfo[t - 1].request_index;
    ffb_beam_per_request_info const &b = bc->beamRequestsInfo[last_slot];
    int const leaves = b.sub_request_num * b.beam_size;
    int const layer = b.current_depth;
    cursor += (bc->tokensInfo[t - 1].abs_depth_in_request - bc->tokensInfo[first].abs_depth_in_request) * b.beam_size;
    Req *r = rm->find(bc->requestsInfo[last_slot].request_guid);
    RM_STATE(r && ssm >= 0 && ssm < rm->d.num_ssms, "beam batch names an unknown request or SSM %d", ssm);
    if (bc->requestsInfo[last_slot].num_tokens_in_batch == 0) {
      continue; // (the reference keeps the previous run's start here too)
    }
    RM_STATE(layer >= 1 && layer < kLayers && leaves >= 0 && leaves <= kBranches,
             "tree layer %d with %d nodes does not fit the speculation tree", layer, leaves);
    RM_STATE(cursor + leaves <= FFB_MAX_NUM_TOKENS * kBranches, "beam result index out of range");
    TreeLayer *tree = &r->spec[(size_t)ssm * kLayers];
    if (layer == 1) { // the root is the last token the request holds
      tree[0].count = 1;
      tree[0].token[0] = r->tokens.back();
      tree[0].prob[0] = 1.f;
      tree[0].parent[0] = -1;
    }
    tree[layer].count = leaves;
    for (int j = 0; j < leaves; j++, cursor++) {
      tree[layer].token[j] = res->token_ids[cursor];
      tree[layer].prob[j] = res->probs[cursor];
      tree[layer].parent[j] = res->parent_id[cursor];
    }
    first = t;
  }
  return FFB_OK;
}

// traverse_beam_tree (request_manager.cc:2743-2815) + merge_dfs_trees for the single-SSM case (:2817-2823): the tree's
// layers 0..max_depth flattened in layer order, depths made absolute.
void flatten_tree(Req &r, int ssm, int max_depth, int root_depth) {
  r.tree_in.clear();
  TreeLayer const *tree = &r.spec[(size_t)ssm * kLayers];
  for (int layer = 0; layer <= max_depth && layer < kLayers; layer++) {
    for (int j = 0; j < tree[layer].count && j < kBranches; j++) {
      r.tree_in.push_back(ffb_tree_node{tree[layer].token[j], layer + root_depth});
    }
  }
}

} // namespace

extern "C" {

int ffb_rm_create(ffb_rm_desc const *desc, ffb_rm_t *out) {
  FFB_REQUIRE(desc && out, "null argument");
  FFB_REQUIRE(desc->max_requests_per_batch > 0 && desc->max_requests_per_batch <= FFB_MAX_NUM_REQUESTS,
              "max_requests_per_batch %d outside 1..%d", desc->max_requests_per_batch, FFB_MAX_NUM_REQUESTS);
  FFB_REQUIRE(desc->max_tokens_per_batch > 0 && desc->max_tokens_per_batch <= FFB_MAX_NUM_TOKENS,
              "max_tokens_per_batch %d outside 1..%d", desc->max_tokens_per_batch, FFB_MAX_NUM_TOKENS);
  FFB_REQUIRE(desc->max_spec_tree_token_num >= 0 && desc->max_spec_tree_token_num <= FFB_MAX_SPEC_TREE_TOKEN_NUM,
              "max_spec_tree_token_num %d outside 0..%d", desc->max_spec_tree_token_num, FFB_MAX_SPEC_TREE_TOKEN_NUM);
  FFB_REQUIRE(desc->max_sequence_length > 0 && desc->num_ssms >= 0 && desc->num_eos_tokens >= 0 &&
                  desc->num_eos_tokens <= 8, "bad max_sequence_length / num_ssms / num_eos_tokens");
  ffb_rm_s *rm = new (std::nothrow) ffb_rm_s();
  FFB_REQUIRE(rm, "out of memory");
  rm->d = *desc;
  *out = rm;
  return FFB_OK;
}

int ffb_rm_destroy(ffb_rm_t rm) {
  delete rm;
  return FFB_OK;
}

int ffb_rm_push_spec_infer_tree_width(ffb_rm_t rm, int tree_width) {
  FFB_REQUIRE(rm && tree_width >= 1 && tree_width <= FFB_MAX_BEAM_WIDTH, "tree width %d outside 1..%d", tree_width,
              FFB_MAX_BEAM_WIDTH);
  rm->widths.push_back(tree_width);
  return FFB_OK;
}

int ffb_rm_register_request(ffb_rm_t rm, int const *prompt_tokens, int num_tokens, int max_length, int max_new_tokens,
                            size_t *guid_out) {
  FFB_REQUIRE(rm && prompt_tokens && num_tokens > 0 && guid_out, "bad argument");
  if (max_length == -1 && max_new_tokens == -1) {
    max_length = rm->d.max_sequence_length - 1;
  }
  if (max_new_tokens != -1) {
    max_length = num_tokens + max_new_tokens;
  }
  FFB_REQUIRE(num_tokens < rm->d.max_sequence_length, "prompt (%d tokens) exceeds max sequence length of %d", num_tokens,
              rm->d.max_sequence_length);
  FFB_REQUIRE(max_length < rm->d.max_sequence_length, "max_length (%d) exceeds max sequence length of %d", max_length,
              rm->d.max_sequence_length);
  Req r;
  r.max_length = max_length;
  r.tokens.assign(prompt_tokens, prompt_tokens + num_tokens);
  r.initial_len = num_tokens;
  r.spec.assign((size_t)rm->d.num_ssms * kLayers, TreeLayer());
  size_t const guid = kFirstGuid + rm->reqs.size();
  rm->reqs.push_back(std::move(r));
  rm->pending.push_back(guid);
  *guid_out = guid;
  return FFB_OK;
}

int ffb_rm_prepare_next_batch(ffb_rm_t rm, ffb_batch_config const *old_bc, ffb_inference_result const *result,
                              ffb_batch_config *new_bc) {
  FFB_REQUIRE(rm && old_bc && result && new_bc, "null argument");
  int const R = rm->d.max_requests_per_batch, cap = rm->d.max_tokens_per_batch;
  FFB_REQUIRE(old_bc->num_tokens >= 0 && old_bc->num_tokens <= FFB_MAX_NUM_TOKENS, "num_tokens out of range");
  // the token each request's newest position produced joins its sequence
  for (int t = 0; t < old_bc->num_tokens; t++) {
    int const slot = old_bc->tokensInfo[t].request_index;
    RM_STATE(slot >= 0 && slot < FFB_MAX_NUM_REQUESTS, "token %d names request slot %d", t, slot);
    Req *r = rm->find(old_bc->requestsInfo[slot].request_guid);
    RM_STATE(r, "token %d belongs to an unknown request", t);
    int const next_pos = old_bc->tokensInfo[t].abs_depth_in_request + 1;
    if (next_pos < (int)r->tokens.size()) {
      continue; // inside the prompt
    }
    RM_STATE(next_pos == (int)r->tokens.size(), "token %d at depth %d is beyond its request (%zu tokens)", t, next_pos - 1,
             r->tokens.size());
    r->tokens.push_back(result->token_ids[t]);
  }
  fresh(new_bc);
  int active = 0, generating = 0;
  for (int i = 0; i < R; i++) {
    if (old_bc->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &o = old_bc->requestsInfo[i];
    Req *r = rm->find(o.request_guid);
    RM_STATE(r && o.num_tokens_in_batch > 0, "slot %d: unknown request or empty slot", i);
    int const done = o.first_token_depth_in_request + o.num_tokens_in_batch;
    RM_STATE(done < (int)r->tokens.size(), "slot %d has processed %d of %zu tokens", i, done, r->tokens.size());
    if (rm->finished_decoding(*r, o.max_length)) {
      if (rm->is_eos(r->tokens.back())) {
        r->tokens.pop_back();
      }
      r->status = FFB_REQ_COMPLETED;
      rm->processed++;
      continue;
    }
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    new_bc->request_completed[i] = false;
    n.first_token_depth_in_request = done;
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = o.request_guid;
    n.peft_model_id = o.peft_model_id;
    std::memcpy(n.peft_model_config_str, o.peft_model_config_str, sizeof(n.peft_model_config_str));
    n.peft_bwd = o.peft_bwd;
    n.max_length = o.max_length;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    if (done + 1 == (int)r->tokens.size()) {
      n.num_tokens_in_batch = 1;
      n.prompt_phase = false;
      generating++;
    } else {
      RM_STATE(o.prompt_phase, "slot %d is behind its sequence but not in the prompt phase", i);
      // leave one position for every later request that will still be decoding
      int reserved = 0;
      for (int k = i + 1; k < R; k++) {
        if (old_bc->request_completed[k]) {
          continue;
        }
        Req *later = rm->find(old_bc->requestsInfo[k].request_guid);
        RM_STATE(later, "slot %d: unknown request", k);
        reserved += rm->finished_decoding(*later, old_bc->requestsInfo[k].max_length) ? 0 : 1;
      }
      n.num_tokens_in_batch = std::min(cap - new_bc->num_tokens - reserved, (int)r->tokens.size() - done);
      n.prompt_phase = true;
    }
    for (int j = 0; j < n.num_tokens_in_batch; j++) {
      RM_STATE(new_bc->num_tokens < FFB_MAX_NUM_TOKENS, "batch overflows MAX_NUM_TOKENS");
      put_token(new_bc, i, done + j, r->tokens[done + j]);
    }
    r->llm_steps++;
  }
  new_bc->num_generation_tokens = generating;
  // free slots take requests from the queue while the batch has room
  for (int i = 0; i < R && !rm->pending.empty(); i++) {
    if (!new_bc->request_completed[i] || new_bc->num_tokens >= cap) {
      continue;
    }
    size_t const guid = rm->pending.front();
    rm->pending.pop_front();
    Req *r = rm->find(guid);
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    n.first_token_depth_in_request = 0;
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = guid;
    n.num_tokens_in_batch = std::min(cap - new_bc->num_tokens, (int)r->tokens.size());
    n.max_length = r->max_length;
    n.peft_model_id = 0;
    n.peft_bwd = false;
    n.prompt_phase = true;
    new_bc->request_completed[i] = false;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    r->llm_steps = 1;
    for (int j = 0; j < n.num_tokens_in_batch; j++) {
      put_token(new_bc, i, j, r->tokens[j]);
    }
    if (new_bc->num_tokens == cap) {
      break;
    }
  }
  return FFB_OK;
}

int ffb_rm_prepare_next_batch_init(ffb_rm_t rm, ffb_tree_verify_batch_config const *old_bc,
                                   ffb_inference_result const *result, int model_id,
                                   ffb_beam_search_batch_config *new_bc) {
  FFB_REQUIRE(rm && old_bc && result && new_bc, "null argument");
  FFB_REQUIRE(rm->d.num_ssms > 0 && model_id >= 0 && model_id < rm->d.num_ssms, "no SSM %d registered", model_id);
  int const R = rm->d.max_requests_per_batch, cap = rm->d.max_tokens_per_batch;
  fresh_beam(new_bc);
  new_bc->model_id = model_id;
  int cursor = 0; // walks the verifier's results in batch order
  int active = 0;
  std::vector<ffb_tree_node> llm_out, accepted;
  std::vector<ffb_commit_entry> kept;
  for (int i = 0; i < R; i++) {
    if (old_bc->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &o = old_bc->requestsInfo[i];
    Req *r = rm->find(o.request_guid);
    RM_STATE(r && old_bc->num_tokens > 0, "slot %d: unknown request or empty verify batch", i);
    // what the verifier said about this request's positions: every prompt position of a request still loading its
    // prompt is a commit candidate; of a running request only the root and the tree above it
    llm_out.clear();
    r->commits.clear();
    r->has_commits = true;
    int const root = (int)r->tokens.size() - 1;
    for (; cursor < old_bc->num_tokens && old_bc->tokensInfo[cursor].request_index == i; cursor++) {
      int const depth = old_bc->tokensInfo[cursor].abs_depth_in_request;
      if (r->status == FFB_REQ_PENDING) {
        r->commits.push_back(ffb_commit_entry{depth, cursor});
      } else if (depth >= root) {
        llm_out.push_back(ffb_tree_node{result->token_ids[cursor], depth + 1});
        r->commits.push_back(ffb_commit_entry{depth, cursor});
      }
    }
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    ffb_beam_per_request_info &b = new_bc->beamRequestsInfo[i];
    if (r->status == FFB_REQ_RUNNING) {
      RM_STATE(!llm_out.empty() && r->tree_in.size() >= llm_out.size() && r->commits.size() >= llm_out.size(),
               "slot %d: verify batch and speculated tree disagree (%zu tree nodes, %zu results)", i, r->tree_in.size(),
               llm_out.size());
      accepted.resize(r->tree_in.size() + 1);
      kept.resize(r->tree_in.size() + 1);
      int n_acc = 0, n_kept = 0;
      int const rc = ffb_traverse_verify_tree(r->tree_in.data(), (int)r->tree_in.size(), llm_out.data(), (int)llm_out.size(),
                                              r->commits.data(), accepted.data(), &n_acc, kept.data(), &n_kept);
      if (rc != FFB_OK) {
        return rc;
      }
      r->commits.assign(kept.begin(), kept.begin() + n_kept);
      if (n_acc + (int)r->tokens.size() >= r->max_length) { // finished: keep what fits
        for (int j = 0; j < n_acc; j++) {
          if (accepted[j].depth < r->max_length) {
            r->tokens.push_back(accepted[j].token_id);
          }
        }
        r->status = FFB_REQ_COMPLETED;
        new_bc->request_completed[i] = true;
        new_bc->request_running[i] = false;
        rm->processed++;
        r->ssm_steps = 0;
        r->tree_in.clear();
        continue;
      }
      // the accepted tokens become the SSM's next "prompt": it catches up on them and proposes from the last one
      new_bc->request_completed[i] = false;
      new_bc->request_running[i] = true;
      n.first_token_depth_in_request = accepted[0].depth;
      n.first_token_offset_in_batch = new_bc->num_tokens;
      n.request_guid = o.request_guid;
      n.max_length = o.max_length;
      n.num_tokens_in_batch = n_acc;
      n.prompt_phase = true;
      new_bc->requestsInfo[active++].batch_config_request_id = i;
      r->ssm_steps = 0;
      root_beam_state(b, rm->width_at(0), std::min(n.max_length - n.first_token_depth_in_request - n_acc, FFB_MAX_BEAM_DEPTH));
      new_bc->sub_requests[i] = 1;
      ffb_bitmask_update(&new_bc->causalMask[i], n_acc, (int)r->tokens.size());
      for (int j = 0; j < n_acc; j++) {
        new_bc->beamTokenInfo[new_bc->num_tokens].sub_request_index = 0;
        put_token(new_bc, i, accepted[j].depth, accepted[j].token_id);
        r->tokens.push_back(accepted[j].token_id);
        if (new_bc->num_tokens == cap) {
          break;
        }
      }
    } else if (r->status == FFB_REQ_PENDING) {
      // still loading its prompt into the verifier: the SSM (which has all of it) idles this round
      RM_STATE(r->ssm_cache_size == r->initial_len, "slot %d: the SSM holds %d of %d prompt tokens", i, r->ssm_cache_size,
               r->initial_len);
      new_bc->request_completed[i] = false;
      new_bc->request_running[i] = false;
      n.first_token_depth_in_request = r->ssm_cache_size;
      n.first_token_offset_in_batch = new_bc->num_tokens;
      n.request_guid = o.request_guid;
      n.max_length = o.max_length;
      n.num_tokens_in_batch = 0;
      new_bc->requestsInfo[active++].batch_config_request_id = i;
      root_beam_state(b, rm->width_at(r->ssm_steps), 0);
      new_bc->sub_requests[i] = 1;
    } else {
      RM_STATE(false, "slot %d holds a request that is neither running nor pending", i);
    }
  }
  // free slots take requests from the queue: their prompt goes to the SSM first
  for (int i = 0; i < R; i++) {
    if (!new_bc->request_completed[i] || rm->pending.empty() || new_bc->num_tokens >= cap) {
      continue;
    }
    size_t const guid = rm->pending.front();
    rm->pending.pop_front();
    Req *r = rm->find(guid);
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    n.first_token_depth_in_request = 0;
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = guid;
    n.num_tokens_in_batch = std::min(cap - new_bc->num_tokens, (int)r->tokens.size());
    n.max_length = r->max_length;
    n.prompt_phase = true;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    r->llm_steps = 0;
    r->ssm_steps = 0;
    root_beam_state(new_bc->beamRequestsInfo[i], rm->width_at(0),
                    std::min(FFB_MAX_BEAM_DEPTH, cap - n.num_tokens_in_batch - 1));
    new_bc->request_completed[i] = false;
    new_bc->request_running[i] = false;
    new_bc->sub_requests[i] = 1;
    for (int j = 0; j < n.num_tokens_in_batch; j++) {
      new_bc->beamTokenInfo[new_bc->num_tokens].sub_request_index = 0;
      put_token(new_bc, i, j, r->tokens[j]);
    }
    ffb_bitmask_init(&new_bc->causalMask[i], n.num_tokens_in_batch);
    r->status = FFB_REQ_PENDING;
    r->ssm_cache_size = n.num_tokens_in_batch;
    if (new_bc->num_tokens == cap) {
      break;
    }
  }
  new_bc->num_generation_tokens = 0;
  return FFB_OK;
}

int ffb_rm_prepare_next_batch_beam(ffb_rm_t rm, ffb_beam_search_batch_config const *old_bc,
                                   ffb_beam_inference_result const *result, ffb_beam_search_batch_config *new_bc) {
  FFB_REQUIRE(rm && old_bc && result && new_bc, "null argument");
  FFB_REQUIRE(old_bc->model_id >= 0 && old_bc->model_id < rm->d.num_ssms, "no SSM %d registered", old_bc->model_id);
  int const R = rm->d.max_requests_per_batch, cap = rm->d.max_tokens_per_batch, ssm = old_bc->model_id;
  int rc = record_ssm_step(rm, old_bc, result);
  if (rc != FFB_OK) {
    return rc;
  }
  fresh_beam(new_bc);
  new_bc->model_id = ssm;
  int active = 0, generating = 0;
  // speculating requests first: one token per node of the layer just filed
  for (int i = 0; i < R; i++) {
    if (old_bc->request_completed[i] || !old_bc->request_running[i]) {
      continue;
    }
    ffb_per_request_info const &o = old_bc->requestsInfo[i];
    ffb_beam_per_request_info const &ob = old_bc->beamRequestsInfo[i];
    Req *r = rm->find(o.request_guid);
    RM_STATE(r && r->status == FFB_REQ_RUNNING, "slot %d: a speculating request must be running", i);
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    ffb_beam_per_request_info &b = new_bc->beamRequestsInfo[i];
    new_bc->request_completed[i] = false;
    new_bc->request_running[i] = true;
    n.first_token_depth_in_request = o.first_token_depth_in_request + o.num_tokens_in_batch;
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = o.request_guid;
    n.max_length = o.max_length;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    r->ssm_steps++;
    b.beam_size = rm->width_at(r->ssm_steps);
    b.max_depth = ob.max_depth;
    b.current_depth = ob.current_depth + 1;
    b.sub_request_num = ob.sub_request_num * ob.beam_size;
    new_bc->sub_requests[i] = old_bc->sub_requests[i] * b.beam_size;
    RM_STATE(b.sub_request_num >= 0 && b.sub_request_num <= kBranches, "slot %d: %d nodes exceed the maximum per tree layer",
             i, b.sub_request_num);
    int const layer = b.current_depth - 1; // update_beam_metadata: the layer record_ssm_step has just written
    RM_STATE(layer >= 0 && layer < kLayers, "slot %d: tree depth %d out of range", i, layer);
    TreeLayer const *tree = &r->spec[(size_t)ssm * kLayers];
    if (b.current_depth != 1) {
      for (int j = 0; j < b.sub_request_num; j++) {
        b.parent_id[j] = tree[layer].parent[j];
        b.probs[j] = tree[layer].prob[j];
        b.tokens[j] = tree[layer].token[j];
      }
    }
    if (n.first_token_depth_in_request >= (int)r->tokens.size()) {
      n.num_tokens_in_batch = 1;
    }
    new_bc->causalMask[i] = old_bc->causalMask[i];
    int widths[kLayers + 1];
    for (int l = 0; l < kLayers; l++) {
      widths[l] = tree[l].count;
    }
    RM_STATE(ob.current_depth <= kLayers, "slot %d: tree depth %d out of range", i, ob.current_depth);
    rc = ffb_bitmask_append(&new_bc->causalMask[i], b.sub_request_num, widths, ob.current_depth);
    if (rc != FFB_OK) {
      return rc == FFB_ERR_INVALID ? FFB_ERR_STATE : rc;
    }
    for (int j = 0; j < n.num_tokens_in_batch; j++) {
      for (int k = 0; k < b.sub_request_num; k++) {
        RM_STATE(new_bc->num_tokens < FFB_MAX_NUM_TOKENS, "batch overflows MAX_NUM_TOKENS");
        new_bc->beamTokenInfo[new_bc->num_tokens].sub_request_index = k;
        put_token(new_bc, i, n.first_token_depth_in_request + j, b.tokens[k]);
        generating++;
      }
    }
  }
  new_bc->speculative_request_num = active;
  // then requests whose prompt the SSM is still loading
  for (int i = 0; i < R; i++) {
    if (old_bc->request_completed[i] || old_bc->request_running[i]) {
      continue;
    }
    ffb_per_request_info const &o = old_bc->requestsInfo[i];
    ffb_beam_per_request_info const &ob = old_bc->beamRequestsInfo[i];
    Req *r = rm->find(o.request_guid);
    RM_STATE(r && r->status == FFB_REQ_PENDING, "slot %d: a request outside the speculation phase must be pending", i);
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    ffb_beam_per_request_info &b = new_bc->beamRequestsInfo[i];
    new_bc->request_completed[i] = false;
    new_bc->request_running[i] = false;
    n.first_token_depth_in_request = o.first_token_depth_in_request + o.num_tokens_in_batch;
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = o.request_guid;
    n.max_length = o.max_length;
    n.prompt_phase = true;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    b.beam_size = 1;
    b.max_depth = ob.max_depth;
    b.current_depth = ob.current_depth;
    b.sub_request_num = ob.sub_request_num;
    new_bc->sub_requests[i] = 1;
    RM_STATE(b.sub_request_num >= 0 && b.sub_request_num <= kBranches, "slot %d: %d nodes exceed the maximum per tree layer",
             i, b.sub_request_num);
    new_bc->causalMask[i] = old_bc->causalMask[i];
    int const left = (int)r->tokens.size() - n.first_token_depth_in_request;
    if (left <= 0) { // the whole prompt is in the SSM's cache: nothing to run for this request
      n.num_tokens_in_batch = 0;
      new_bc->causalMask[i].this_layer_size = 0;
      b.sub_request_num = 0;
    } else {
      n.num_tokens_in_batch = std::min(cap - new_bc->num_tokens - R + i, left);
      r->ssm_cache_size += n.num_tokens_in_batch;
      ffb_bitmask_append_pending(&new_bc->causalMask[i], n.num_tokens_in_batch);
    }
    for (int j = 0; j < n.num_tokens_in_batch; j++) {
      for (int k = 0; k < b.sub_request_num; k++) {
        RM_STATE(new_bc->num_tokens < FFB_MAX_NUM_TOKENS, "batch overflows MAX_NUM_TOKENS");
        new_bc->beamTokenInfo[new_bc->num_tokens].sub_request_index = k;
        put_token(new_bc, i, n.first_token_depth_in_request + j,
                  r->tokens[r->tokens.size() - n.num_tokens_in_batch + j]);
      }
    }
  }
  new_bc->num_generation_tokens = generating;
  return FFB_OK;
}

int ffb_rm_prepare_next_batch_verify(ffb_rm_t rm, ffb_beam_search_batch_config const *const *old_batches,
                                     int num_batches, ffb_tree_verify_batch_config *new_bc) {
  FFB_REQUIRE(rm && old_batches && new_bc, "null argument");
  FFB_REQUIRE(num_batches == 1 && old_batches[0], "one SSM is supported (the reference's merge_dfs_trees asserts the same)");
  ffb_beam_search_batch_config const *last = old_batches[0];
  FFB_REQUIRE(last->model_id >= 0 && last->model_id < rm->d.num_ssms, "no SSM %d registered", last->model_id);
  int const R = rm->d.max_requests_per_batch, cap = rm->max_verify_tokens();
  fresh(new_bc);
  new_bc->num_tokens_to_commit = 0;
  // room left for prompts once every speculating request got its worst-case tree and every loading one a token
  int prompt_room = cap;
  for (int i = 0; i < R; i++) {
    if (!last->request_completed[i]) {
      prompt_room -= last->request_running[i] ? FFB_MAX_BEAM_DEPTH + 1 : 1;
    }
  }
  int active = 0;
  for (int i = 0; i < R; i++) {
    if (last->request_completed[i]) {
      continue;
    }
    ffb_per_request_info const &o = last->requestsInfo[i];
    Req *r = rm->find(o.request_guid);
    RM_STATE(r, "slot %d: unknown request", i);
    r->llm_steps++;
    ffb_per_request_info &n = new_bc->requestsInfo[i];
    n.first_token_offset_in_batch = new_bc->num_tokens;
    n.request_guid = o.request_guid;
    n.max_length = o.max_length;
    new_bc->requestsInfo[active++].batch_config_request_id = i;
    new_bc->causalMask[i] = last->causalMask[i];
    new_bc->request_completed[i] = false;
    // tokens accepted (or prompt positions loaded) in the previous verify batch move to their final cache rows
    if (r->has_commits) {
      for (ffb_commit_entry const &c : r->commits) {
        RM_STATE(new_bc->num_tokens_to_commit < FFB_MAX_NUM_TOKENS, "commit list overflows MAX_NUM_TOKENS");
        ffb_committed_token_info &ct = new_bc->committed_tokens[new_bc->num_tokens_to_commit++];
        ct.token_index = c.index_in_batch;
        ct.request_index = i;
        ct.token_depth = c.depth;
        r->llm_cache_size++;
      }
    }
    if (r->status == FFB_REQ_RUNNING) {
      new_bc->request_running[i] = true;
      int const root = (int)r->tokens.size() - 1;
      flatten_tree(*r, last->model_id, last->beamRequestsInfo[i].max_depth, root);
      RM_STATE(!r->tree_in.empty(), "slot %d: empty speculation tree", i);
      n.first_token_depth_in_request = root;
      n.num_tokens_in_batch = 0;
      RM_STATE(new_bc->num_tokens < cap, "exceeding the space available in the TreeVerify batch");
      put_token(new_bc, i, root, r->tokens.back());
      n.num_tokens_in_batch++;
      bool cut = false;
      for (size_t j = 1; j < r->tree_in.size(); j++) {
        put_token(new_bc, i, r->tree_in[j].depth, r->tree_in[j].token_id);
        n.num_tokens_in_batch++;
        if (new_bc->num_tokens == cap && j != r->tree_in.size() - 1) {
          cut = true;
          break;
        }
      }
      if (cut) { // drop the layer the batch boundary cut in two
        for (int j = new_bc->num_tokens - 1; j >= 1; j--) {
          new_bc->num_tokens--;
          n.num_tokens_in_batch--;
          if (new_bc->tokensInfo[j].abs_depth_in_request != new_bc->tokensInfo[j - 1].abs_depth_in_request) {
            break;
          }
        }
      }
    } else if (r->status == FFB_REQ_PENDING) {
      new_bc->request_running[i] = false;
      n.first_token_depth_in_request = r->llm_cache_size;
      n.num_tokens_in_batch = std::min(prompt_room, r->initial_len - n.first_token_depth_in_request);
      prompt_room -= n.num_tokens_in_batch;
      bool launch = false;
      if (r->llm_cache_size < r->initial_len) {
        for (int j = 0; j < n.num_tokens_in_batch; j++) {
          RM_STATE(new_bc->num_tokens < cap, "exceeding the space available in the TreeVerify batch");
          put_token(new_bc, i, r->llm_cache_size + j, r->tokens[r->llm_cache_size + j]);
        }
        launch = n.num_tokens_in_batch + r->llm_cache_size >= r->initial_len;
      } else if (cap - new_bc->num_tokens > 0) {
        put_token(new_bc, i, (int)r->tokens.size() - 1, r->tokens.back());
        n.num_tokens_in_batch++;
        launch = true;
      }
      if (launch) { // the whole prompt is (about to be) in the verifier's cache: speculation starts next round
        r->status = FFB_REQ_RUNNING;
        new_bc->request_running[i] = true;
        n.prompt_phase = true;
        r->tree_in.assign(1, ffb_tree_node{r->tokens.back(), (int)r->tokens.size() - 1});
      }
    } else {
      RM_STATE(false, "slot %d holds a request that is neither running nor pending", i);
    }
  }
  return FFB_OK;
}

int ffb_rm_request_status(ffb_rm_t rm, size_t guid, int *status_out) {
  Req *r = rm ? rm->find(guid) : nullptr;
  FFB_REQUIRE(r && status_out, "unknown request");
  *status_out = r->status;
  return FFB_OK;
}

int ffb_rm_request_tokens(ffb_rm_t rm, size_t guid, int *tokens_out, int capacity, int *num_tokens) {
  Req *r = rm ? rm->find(guid) : nullptr;
  FFB_REQUIRE(r && num_tokens && (tokens_out || capacity <= 0), "unknown request");
  *num_tokens = (int)r->tokens.size();
  for (int i = 0; i < capacity && i < (int)r->tokens.size(); i++) {
    tokens_out[i] = r->tokens[i];
  }
  return FFB_OK;
}

int ffb_rm_request_steps(ffb_rm_t rm, size_t guid, int *llm_decoding_steps, int *ssm_decoding_steps) {
  Req *r = rm ? rm->find(guid) : nullptr;
  FFB_REQUIRE(r, "unknown request");
  if (llm_decoding_steps) {
    *llm_decoding_steps = r->llm_steps;
  }
  if (ssm_decoding_steps) {
    *ssm_decoding_steps = r->ssm_steps;
  }
  return FFB_OK;
}

int ffb_rm_num_pending(ffb_rm_t rm) { return rm ? (int)rm->pending.size() : 0; }
size_t ffb_rm_num_processed_requests(ffb_rm_t rm) { return rm ? rm->processed : 0; }

int ffb_fetch_inference_result(int const *indices_dev, int num_tokens, ffb_inference_result *result_host,
                               ffb_stream_t stream) {
  FFB_REQUIRE(indices_dev && result_host && num_tokens >= 0 && num_tokens <= FFB_MAX_NUM_TOKENS, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  FFB_CUDA_TRY(cudaMemcpyAsync(result_host->token_ids, indices_dev, sizeof(int) * (size_t)num_tokens,
                               cudaMemcpyDeviceToHost, st));
  FFB_CUDA_TRY(cudaStreamSynchronize(st));
  return FFB_OK;
}

int ffb_fetch_beam_inference_result(int const *indices_dev, float const *probs_dev, int num_tokens, int k,
                                    ffb_beam_inference_result *result_host, ffb_stream_t stream) {
  FFB_REQUIRE(indices_dev && probs_dev && result_host && num_tokens >= 0 && k >= 1 && k <= kBranches &&
                  num_tokens <= FFB_MAX_NUM_TOKENS, "bad argument");
  cudaStream_t st = (cudaStream_t)stream;
  size_t const n = (size_t)num_tokens * k;
  FFB_CUDA_TRY(cudaMemcpyAsync(result_host->token_ids, indices_dev, sizeof(int) * n, cudaMemcpyDeviceToHost, st));
  FFB_CUDA_TRY(cudaMemcpyAsync(result_host->probs, probs_dev, sizeof(float) * n, cudaMemcpyDeviceToHost, st));
  std::memset(result_host->parent_id, 0, sizeof(int) * n);
  FFB_CUDA_TRY(cudaStreamSynchronize(st));
  return FFB_OK;
}

} // extern "C"
