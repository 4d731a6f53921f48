// ref_host_wrap.cpp — compiles the REFERENCE's own host-side serving control (SpecInfer tree logic + batch builders) into
// oracle/_ref/libref_host.so.  TEST INFRASTRUCTURE ONLY: tests/test_request_manager.py and tests/test_host_tree.py run
// the product's ffb_rm_* / ffb_bitmask_* / ffb_traverse_verify_tree and the oracle's restatement beside it and compare
// every batch descriptor field for field.  Nothing is copied into this repository: oracle/Makefile streams, with
// oracle/lift_fn.awk, the text of
//     struct Request, struct BeamTree, class RequestManager          include/flexflow/request_manager.h:64-357
//     RequestManager::RequestManager() and the max_* setters/getters  src/runtime/request_manager.cc:102-171
//     is_eos_token, check_inf_req_completion                          :636-657
//     prepare_next_batch (incremental decoding)                       :713-1138
//     prepare_next_batch_init / _beam / _verify                       :1170-1580, :1610-1892, :1923-2215
//     store_beam_metadata, update_beam_metadata                       :2217-2379
//     initBitMask, updateBitMask, appendPendingRequest, appendBitMask :2382-2517
//     traverse_verify_tree, traverse_beam_tree, merge_dfs_trees       :2583-2888
//     BatchConfig / BeamSearchBatchConfig / TreeVerifyBatchConfig constructors and max_*() statics
//                                                                     src/runtime/batch_config.cc:57-150 etc.
// into temporary files that are included below and deleted after the compile.  The reference's own headers
// flexflow/inference.h -> batch_config.h, ffconst.h, fftype.h are read where they lie (-I $(REF)/include), legion.h is
// the stand-in of oracle/ref_stubs.
//
// Supplied here instead of the files' surroundings (Legion, tokenizers_cpp, nlohmann::json, PEFT): forward declarations
// of the Legion / model types the class only names, a Tokenizer whose Decode returns "", a logger that discards, a
// clock that returns 0, PEFTModelID's trivial members (src/runtime/fftype.cc:39-62), a LoraLinearConfig with an empty
// serialisation, and the C entry points at the bottom.  The serving loop itself (serve_spec_infer,
// request_manager.cc:3083-3173: init -> MAX_BEAM_DEPTH x (SSM, beam) -> verify -> LLM) is driven from the tests.
#include <cassert>
#include <cstdarg>
#include <cstring>
#include <fstream>
#include <future>
#include <iomanip>
#include <iostream>
#include <mutex>
#include <queue>
#include <set>
#include <sstream>
#include <stack>
#include <string>
#include <unistd.h>
#include <unordered_map>
#include <vector>

#include "flexflow/inference.h"

namespace Legion {
struct Task;
struct PhysicalRegion;
struct Runtime;
struct FutureMap;
typedef void *Context;
} // namespace Legion
namespace Realm {
struct Clock {
  static double current_time_in_microseconds() { return 0.0; }
};
} // namespace Realm
namespace tokenizers {
struct Tokenizer {
  std::string Decode(std::vector<int> const &) { return std::string(); }
  std::vector<int> Encode(std::string const &) { return {}; }
};
} // namespace tokenizers

namespace FlexFlow {

using Legion::Context;
using Legion::Runtime;
using Legion::Task;
using Legion::PhysicalRegion;
using std::min;
class FFModel;
class FileDataLoader;
struct ParallelTensorBase;
typedef ParallelTensorBase *ParallelTensor;
typedef Legion::Future BeamInferenceResultFuture_unused;

struct LoraLinearConfig {
  std::string serialize_to_json_string(int = -1) const { return std::string(); }
};

// src/runtime/fftype.cc:39-62 (the .cc includes flexflow/config.h, which needs Legion)
const PEFTModelID PEFTModelID::NO_ID = PEFTModelID();
PEFTModelID::PEFTModelID() : id(0) {}
PEFTModelID::PEFTModelID(size_t _id) : id(_id) {}
bool operator==(PEFTModelID const &lhs, PEFTModelID const &rhs) { return lhs.id == rhs.id; }
bool operator!=(PEFTModelID const &lhs, PEFTModelID const &rhs) { return !(lhs == rhs); }
std::ostream &operator<<(std::ostream &os, PEFTModelID const &p) { return os << p.id; }

struct DiscardLogger {
  struct Sink {
    template <typename T>
    Sink &operator<<(T const &) { return *this; }
  };
  void print(char const *, ...) {}
  Sink debug() { return Sink(); }
};
static DiscardLogger log_req_mgr;

using tokenizers::Tokenizer;

#define private public
#include REF_RM_CLASS_INC
#undef private

// ---- definitions the lifted text calls but which are not lifted (trivial or Legion-bound) ----
bool RequestManager::inference_finished = false;
static RequestManager *request_manager_singleton = nullptr;
RequestManager *RequestManager::get_request_manager() {
  if (request_manager_singleton == nullptr) {
    request_manager_singleton = new RequestManager();
  }
  return request_manager_singleton;
}
void RequestManager::trigger_request_completion_future(RequestGuid const &) {}
size_t RequestManager::get_num_ssms() { return ssm_models.size(); }
int RequestManager::get_max_concurrent_adapters() { return max_concurrent_adapters; }
LoraLinearConfig const &RequestManager::get_peft_config(PEFTModelID const &id) { return peft_configs[id]; }
void RequestManager::add_peft_config_to_request_info(BatchConfig &, int, LoraLinearConfig const &) {}
void BatchConfig::print() const {}
void BeamSearchBatchConfig::print() const {}
void TreeVerifyBatchConfig::print() const {}
std::ostream &operator<<(std::ostream &os, BatchConfig const &) { return os; }
BeamSearchBatchConfig::~BeamSearchBatchConfig() {}
TreeVerifyBatchConfig::~TreeVerifyBatchConfig() {}
InferenceMode BatchConfig::get_mode() const { return INC_DECODING_MODE; }
InferenceMode BeamSearchBatchConfig::get_mode() const { return BEAM_SEARCH_MODE; }
InferenceMode TreeVerifyBatchConfig::get_mode() const { return TREE_VERIFY_MODE; }

#include REF_RM_FUNCS_INC

} // namespace FlexFlow

using namespace FlexFlow;

namespace {
// the lifted code prints progress to std::cout / printf; keep the test output clean
struct Quiet {
  std::ostringstream sink;
  std::streambuf *old;
  int saved_fd;
  Quiet() : old(std::cout.rdbuf(sink.rdbuf())) {
    fflush(stdout);
    saved_fd = dup(1);
    int nul = open_null();
    dup2(nul, 1);
    close(nul);
  }
  ~Quiet() {
    std::cout.rdbuf(old);
    fflush(stdout);
    dup2(saved_fd, 1);
    close(saved_fd);
  }
  static int open_null();
};
} // namespace
#include <fcntl.h>
int Quiet::open_null() { return open("/dev/null", O_WRONLY); }

namespace {
template <typename BC>
void load_bc(BC &dst, void const *src) { // keep dst's vptr, take the payload
  std::memcpy((char *)&dst + sizeof(void *), (char const *)src + sizeof(void *), sizeof(BC) - sizeof(void *));
}
template <typename BC>
void store_bc(void *dst, BC const &src) {
  std::memcpy(dst, (void const *)&src, sizeof(BC));
}
RequestManager *rm() { return RequestManager::get_request_manager(); }
} // namespace

extern "C" {

// One RequestManager per process (BatchConfig::max_requests_per_batch() reads the singleton): ref_rm_reset replaces it.
void ref_rm_reset(int max_requests, int max_tokens, int max_spec_tree_tokens, int max_seq_len, int num_ssms,
                  int const *tree_widths, int n_widths, int const *eos, int n_eos) {
  delete request_manager_singleton;
  request_manager_singleton = new RequestManager();
  RequestManager *r = request_manager_singleton;
  r->set_max_requests_per_batch(max_requests);
  r->set_max_tokens_per_batch(max_tokens);
  r->set_max_spec_tree_token_num(max_spec_tree_tokens);
  r->set_max_sequence_length(max_seq_len);
  for (int i = 0; i < n_widths; i++) {
    r->push_spec_infer_tree_width(tree_widths[i]);
  }
  for (int i = 0; i < n_eos; i++) {
    r->eos_token_ids.push_back(eos[i]);
  }
  r->ssm_models.assign(num_ssms, nullptr);
  r->tokenizer_.reset(new Tokenizer());
  r->model_type = ModelType::LLAMA;
  r->bos_token_id = -1;
}

// register_new_request (request_manager.cc:334-441) minus the tokenizer: the caller hands the token ids.  The fields
// set are the ones that function sets for an inference request.
size_t ref_rm_register_request(int const *tokens, int n, int max_length) {
  RequestManager *r = rm();
  Request request;
  request.status = Request::PENDING;
  request.guid = r->next_available_guid++;
  request.max_length = max_length;
  request.tokens.assign(tokens, tokens + n);
  request.initial_len = n;
  for (size_t i = 0; i < r->get_num_ssms(); i++) {
    request.beam_trees.push_back(BeamTree{});
  }
  r->pending_infr_request_queue.push(request);
  r->all_requests[request.guid] = request;
  GenerationResult gr;
  gr.guid = request.guid;
  gr.input_tokens = request.tokens;
  gr.output_tokens = request.tokens;
  r->request_generation_results[request.guid] = gr;
  r->profiling_requests[request.guid] = RequestManager::ProfileInfo();
  return request.guid;
}

void ref_rm_prepare_next_batch(void const *old_bc, void const *result, void *new_bc) {
  Quiet q;
  BatchConfig in;
  load_bc(in, old_bc);
  BatchConfig out = rm()->prepare_next_batch(in, *(InferenceResult const *)result);
  store_bc(new_bc, out);
}
void ref_rm_prepare_next_batch_init(void const *old_tree_bc, void const *result, int model_id, void *new_beam_bc) {
  Quiet q;
  TreeVerifyBatchConfig in;
  load_bc(in, old_tree_bc);
  BeamSearchBatchConfig out = rm()->prepare_next_batch_init(in, *(InferenceResult const *)result, model_id);
  store_bc(new_beam_bc, out);
}
void ref_rm_prepare_next_batch_beam(void const *old_beam_bc, void const *beam_result, void *new_beam_bc) {
  Quiet q;
  BeamSearchBatchConfig in;
  load_bc(in, old_beam_bc);
  BeamSearchBatchConfig out = rm()->prepare_next_batch_beam(in, *(BeamInferenceResult const *)beam_result);
  store_bc(new_beam_bc, out);
}
void ref_rm_prepare_next_batch_verify(void const *const *old_beam_bcs, int n, void *new_tree_bc) {
  Quiet q;
  std::vector<BeamSearchBatchConfig> olds(n);
  for (int i = 0; i < n; i++) {
    load_bc(olds[i], old_beam_bcs[i]);
  }
  TreeVerifyBatchConfig out = rm()->prepare_next_batch_verify(olds);
  store_bc(new_tree_bc, out);
}

int ref_rm_request_tokens(size_t guid, int *out, int cap) {
  auto &t = rm()->all_requests[guid].tokens;
  int n = (int)t.size();
  for (int i = 0; i < n && i < cap; i++) {
    out[i] = t[i];
  }
  return n;
}
int ref_rm_request_status(size_t guid) { return (int)rm()->all_requests[guid].status; }
size_t ref_rm_num_processed() { return rm()->num_processed_requests; }
int ref_rm_pending() { return (int)rm()->pending_infr_request_queue.size(); }
int ref_rm_decoding_steps(size_t guid, int *ssm_steps) {
  auto &p = rm()->profiling_requests[guid];
  if (ssm_steps) {
    *ssm_steps = p.ssm_decoding_steps;
  }
  return p.llm_decoding_steps;
}

size_t ref_sizeof(int what) {
  switch (what) {
    case 0: return sizeof(BatchConfig);
    case 1: return sizeof(BeamSearchBatchConfig);
    case 2: return sizeof(TreeVerifyBatchConfig);
    case 3: return sizeof(InferenceResult);
    case 4: return sizeof(BeamInferenceResult);
    default: return 0;
  }
}

// ---- the bit-mask functions and the accept walk on their own ----
void ref_init_bitmask(void *bm, int init_length) { rm()->initBitMask(*(BatchConfig::BitMask *)bm, init_length); }
void ref_update_bitmask(void *bm, int init_length, int non_tree_size) {
  rm()->updateBitMask(*(BatchConfig::BitMask *)bm, init_length, non_tree_size);
}
void ref_append_pending(void *bm, int init_length) { rm()->appendPendingRequest(*(BatchConfig::BitMask *)bm, init_length); }
// layer_nodes[i] = treeLayers[i].nodes_num_this_layer for i < n_layers
void ref_append_bitmask(void *bm, int new_nodes, int pre_beam_size, int old_sub_num, int const *layer_nodes, int n_layers,
                        int current_depth) {
  BeamTree tree{};
  for (int i = 0; i < n_layers && i <= BeamSearchBatchConfig::MAX_BEAM_DEPTH; i++) {
    tree.treeLayers[i].nodes_num_this_layer = layer_nodes[i];
  }
  rm()->appendBitMask(*(BatchConfig::BitMask *)bm, new_nodes, pre_beam_size, old_sub_num, tree, current_depth);
}
// in_tree / out_tree: (token, depth) pairs; committed_in: (depth, index) pairs of committed_tokens[guid].
// Returns the number of verified pairs; *n_committed the length of the new committed_tokens[guid].
int ref_traverse_verify_tree(int const *in_tree, int n_in, int const *out_tree, int n_out, int const *committed_in,
                             int n_commit_in, int *verified, int *committed_out, int *n_committed) {
  Quiet q;
  size_t const guid = 4242;
  std::vector<std::pair<BatchConfig::TokenId, int>> a, b;
  for (int i = 0; i < n_in; i++) {
    a.emplace_back(in_tree[2 * i], in_tree[2 * i + 1]);
  }
  for (int i = 0; i < n_out; i++) {
    b.emplace_back(out_tree[2 * i], out_tree[2 * i + 1]);
  }
  auto &ct = rm()->committed_tokens[guid];
  ct.clear();
  for (int i = 0; i < n_commit_in; i++) {
    ct.emplace_back(committed_in[2 * i], committed_in[2 * i + 1]);
  }
  auto v = rm()->traverse_verify_tree(guid, a, b);
  for (size_t i = 0; i < v.size(); i++) {
    verified[2 * i] = v[i].first;
    verified[2 * i + 1] = v[i].second;
  }
  auto &nc = rm()->committed_tokens[guid];
  for (size_t i = 0; i < nc.size(); i++) {
    committed_out[2 * i] = nc[i].first;
    committed_out[2 * i + 1] = nc[i].second;
  }
  *n_committed = (int)nc.size();
  return (int)v.size();
}

} // extern "C"
