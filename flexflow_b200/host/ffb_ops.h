// ffb_ops.h — C++ host-side mirror of the reference's operator interface for the attention hot path,
// header-only on top of the C ABI (include/ffb200.h).  Names, argument order and meaning follow the
// reference so that its op tasks can call these instead of their own wrappers:
//   IncMultiHeadSelfAttention::inference_kernel_wrapper     include/flexflow/ops/inc_multihead_self_attention.h:115-119
//   TreeIncMultiHeadSelfAttention::inference_kernel_wrapper include/flexflow/ops/tree_inc_multihead_self_attention.h:104-108
//   SpecIncMultiHeadSelfAttention::inference_kernel_wrapper include/flexflow/ops/spec_inc_multihead_self_attention.h:101-106
//   Kernels::Linear::{forward,inference}_kernel_wrapper     src/ops/kernels/linear_kernels.cu:133-140,187-195
//   Kernels::AllReduce::inference_kernel_wrapper            include/flexflow/parallel_ops/kernels/allreduce_kernels.h:28-31
// Differences: errors throw ffb::Error instead of assert()/exit(); the CUDA stream is passed explicitly
// (the reference pulls it from Legion, src/runtime/cuda_helper.cu:18-23).
#pragma once
#include <stdexcept>
#include <string>
#include <vector>

#include "../../include/ffb200.h"

namespace ffb {

struct Error : std::runtime_error {
  int code;
  Error(int c, std::string const &what) : std::runtime_error(what), code(c) {}
};
inline void check(int rc, char const *what) {
  if (rc != FFB_OK) {
    throw Error(rc, std::string(what) + ": " + ffb_last_error());
  }
}

// include/flexflow/accessor.h:55-87 — {data_type, domain, ptr}; the domain is reduced to what the path reads
struct GenericTensorAccessorR {
  int data_type;
  void const *ptr;
  size_t hidden_dim; // innermost extent (domain.hi()[0]-domain.lo()[0]+1)
};
struct GenericTensorAccessorW {
  int data_type;
  void *ptr;
  size_t hidden_dim;
};

// include/flexflow/inference.h:46-66
struct RotaryEmbeddingMeta {
  bool apply_rotary_embedding = false;
  float rope_theta = 10000.0f;
  std::string rope_type = "default";
  float factor = 8.0f, low_freq_factor = 1.0f, high_freq_factor = 4.0f;
  int original_max_position_embeddings = 8192;
};

// The slice of FFHandler the path uses (include/flexflow/config.h:91-113)
struct FFHandler {
  ffb_combined_batch_config_meta *batch_config_metadata = nullptr; // device
  void *staging_pinned = nullptr;                                  // host, pinned, same size
  ffb_comm_t ncclComm = nullptr;
  // RequestManager::load_batch_config_task, src/runtime/request_manager.cu:83-165
  void load_batch_config(void const *bc, int mode, ffb_stream_t stream) const {
    check(ffb_load_batch_config(bc, mode, staging_pinned, batch_config_metadata, stream), "ffb_load_batch_config");
  }
};

// IncMultiHeadSelfAttentionMeta (inc_multihead_self_attention.h:139-202; ctor inc_…cu:1621-1807).
// _num_q_heads/_num_kv_heads are the per-shard counts (inc_…cc:516-519).
class IncMultiHeadSelfAttentionMeta {
public:
  IncMultiHeadSelfAttentionMeta(FFHandler handler, int infer_mode, int data_type, int head_dim,
                                RotaryEmbeddingMeta const &rope, bool scaling_query, bool qk_prod_scaling,
                                bool position_bias, float scaling_factor, int global_num_q_heads, int _num_q_heads,
                                int _num_kv_heads, int max_requests_per_batch, int max_tokens_per_batch,
                                int max_sequence_length, int max_spec_tree_token_num)
      : handle(handler) {
    ffb_attn_desc d{};
    d.mode = infer_mode; d.dtype = data_type; d.num_q_heads = _num_q_heads; d.num_kv_heads = _num_kv_heads;
    d.head_dim = head_dim; d.global_num_q_heads = global_num_q_heads;
    d.apply_rotary_embedding = rope.apply_rotary_embedding; d.rope_theta = rope.rope_theta;
    d.rope_llama3 = rope.rope_type == "llama3"; d.rope_factor = rope.factor;
    d.rope_low_freq_factor = rope.low_freq_factor; d.rope_high_freq_factor = rope.high_freq_factor;
    d.rope_original_max_position_embeddings = rope.original_max_position_embeddings;
    d.scaling_query = scaling_query; d.scaling_factor = scaling_factor; d.qk_prod_scaling = qk_prod_scaling;
    d.position_bias = position_bias; d.max_requests = max_requests_per_batch;
    d.max_sequence_length = max_sequence_length;
    d.max_spec_tree_tokens = infer_mode == FFB_INC_DECODING_MODE ? 0 : max_spec_tree_token_num;
    d.max_tokens_per_batch = max_tokens_per_batch;
    check(ffb_attn_create(&d, &h_), "ffb_attn_create");
  }
  ~IncMultiHeadSelfAttentionMeta() { ffb_attn_destroy(h_); }
  IncMultiHeadSelfAttentionMeta(IncMultiHeadSelfAttentionMeta const &) = delete;
  IncMultiHeadSelfAttentionMeta &operator=(IncMultiHeadSelfAttentionMeta const &) = delete;
  FFHandler handle;
  bool profiling = false;
  ffb_attn_t raw() const { return h_; }

private:
  ffb_attn_t h_ = nullptr;
};
using TreeIncMultiHeadSelfAttentionMeta = IncMultiHeadSelfAttentionMeta; // mode = FFB_TREE_VERIFY_MODE
using SpecIncMultiHeadSelfAttentionMeta = IncMultiHeadSelfAttentionMeta; // mode = FFB_BEAM_SEARCH_MODE

struct IncMultiHeadSelfAttention {
  static void inference_kernel_wrapper(IncMultiHeadSelfAttentionMeta *m, ffb_batch_config const *bc, int shard_id,
                                       GenericTensorAccessorR const &input, GenericTensorAccessorW const &output,
                                       ffb_stream_t stream) {
    if (input.data_type != output.data_type) {
      throw Error(FFB_ERR_INVALID, "input/output data types differ"); // assert(input.data_type == output.data_type)
    }
    check(ffb_inc_attn_forward(m->raw(), bc, m->handle.batch_config_metadata, shard_id, input.ptr, output.ptr, stream),
          "ffb_inc_attn_forward");
  }
};
struct TreeIncMultiHeadSelfAttention {
  static void inference_kernel_wrapper(TreeIncMultiHeadSelfAttentionMeta *m, ffb_tree_verify_batch_config const *bc,
                                       int shard_id, GenericTensorAccessorR const &input,
                                       GenericTensorAccessorW const &output, ffb_stream_t stream) {
    check(ffb_tree_attn_forward(m->raw(), bc, m->handle.batch_config_metadata, shard_id, input.ptr, output.ptr, stream),
          "ffb_tree_attn_forward");
  }
};
struct SpecIncMultiHeadSelfAttention {
  static void inference_kernel_wrapper(SpecIncMultiHeadSelfAttentionMeta const *m,
                                       ffb_beam_search_batch_config const *bc, int shard_id,
                                       GenericTensorAccessorR const &input, GenericTensorAccessorW const &output,
                                       ffb_stream_t stream) {
    check(ffb_spec_attn_forward(m->raw(), bc, m->handle.batch_config_metadata, shard_id, input.ptr, output.ptr, stream),
          "ffb_spec_attn_forward");
  }
};

class LinearMeta {
public:
  LinearMeta(FFHandler handler, int data_type, int quantization_type, int in_dim, int out_dim, int max_batch)
      : handle(handler) {
    ffb_linear_desc d{data_type, quantization_type, in_dim, out_dim, max_batch, 0};
    check(ffb_linear_create(&d, &h_), "ffb_linear_create");
  }
  ~LinearMeta() { ffb_linear_destroy(h_); }
  LinearMeta(LinearMeta const &) = delete;
  LinearMeta &operator=(LinearMeta const &) = delete;
  FFHandler handle;
  ffb_linear_t raw() const { return h_; }

private:
  ffb_linear_t h_ = nullptr;
};

namespace Kernels {
namespace Linear {
inline void forward_kernel_wrapper(LinearMeta const *m, void const *input_ptr, void *output_ptr,
                                   void const *filter_ptr, void const *bias_ptr, int in_dim, int out_dim,
                                   int batch_size, ffb_stream_t stream) {
  (void)in_dim;
  (void)out_dim;
  check(ffb_linear_forward(m->raw(), input_ptr, output_ptr, filter_ptr, bias_ptr, batch_size, stream), "ffb_linear_forward");
}
inline void inference_kernel_wrapper(LinearMeta *m, ffb_batch_config const *bc, void const *input_ptr,
                                     void *output_ptr, void const *filter_ptr, void const *bias_ptr, int in_dim,
                                     int out_dim, int batch_size, ffb_stream_t stream) {
  (void)bc;
  forward_kernel_wrapper(m, input_ptr, output_ptr, filter_ptr, bias_ptr, in_dim, out_dim, batch_size, stream);
}
} // namespace Linear
namespace AllReduce {
struct AllReduceMeta {
  FFHandler handle;
};
inline void inference_kernel_wrapper(AllReduceMeta const *m, ffb_batch_config const *bc,
                                     GenericTensorAccessorR const &input, GenericTensorAccessorW const &output,
                                     ffb_stream_t stream) {
  size_t const num_elements = (size_t)bc->num_tokens * input.hidden_dim; // allreduce_kernels.cu:64-65
  check(ffb_allreduce(m->handle.ncclComm, input.ptr, output.ptr, num_elements, input.data_type, stream), "ffb_allreduce");
}
} // namespace AllReduce

// The row operators either side of the chain.
namespace ResidualRMSNorm {
struct ResidualRMSNormMeta { // include/flexflow/ops/residual_rms_norm.h: eps, in_dim, batch_size
  float eps;
  int in_dim;
};
// Kernels::ResidualRMSNorm::inference_kernel_wrapper (src/ops/kernels/residual_rms_norm_kernels.cu:205-232)
inline void inference_kernel_wrapper(ResidualRMSNormMeta *m, ffb_batch_config const *bc,
                                     GenericTensorAccessorR const &input1, GenericTensorAccessorR const &input2,
                                     GenericTensorAccessorR const &weight, GenericTensorAccessorW const &residual_output,
                                     GenericTensorAccessorW const &output, ffb_stream_t stream) {
  check(ffb_residual_rms_norm(input1.ptr, input2.ptr, weight.ptr, residual_output.ptr, output.ptr, bc->num_tokens,
                              m->in_dim, m->eps, output.data_type, stream), "ffb_residual_rms_norm");
}
} // namespace ResidualRMSNorm
namespace RMSNorm {
// Kernels::RMSNorm::inference_kernel_wrapper (src/ops/kernels/rms_norm_kernels.cu:148-190)
inline void inference_kernel_wrapper(ResidualRMSNorm::ResidualRMSNormMeta *m, ffb_batch_config const *bc,
                                     GenericTensorAccessorR const &input, GenericTensorAccessorR const &weight,
                                     GenericTensorAccessorW const &output, ffb_stream_t stream) {
  check(ffb_residual_rms_norm(input.ptr, nullptr, weight.ptr, nullptr, output.ptr, bc->num_tokens, m->in_dim, m->eps,
                              output.data_type, stream), "ffb_residual_rms_norm");
}
} // namespace RMSNorm
} // namespace Kernels

// ArgMax::forward_kernel_wrapper (src/ops/argmax.cu:156-206) / ArgTopK::forward_kernel_wrapper (src/ops/arg_topk.cu:451-541)
struct ArgMax {
  static void forward_kernel_wrapper(bool beam_search, GenericTensorAccessorR const &input, int *indices, float *probs,
                                     int *parent, int batch_size, int length, ffb_stream_t stream) {
    check(ffb_argmax(input.ptr, indices, beam_search ? probs : nullptr, beam_search ? parent : nullptr, batch_size, length,
                     input.data_type, stream), "ffb_argmax");
  }
};
struct ArgTopK {
  static void forward_kernel_wrapper(bool speculative_decoding, GenericTensorAccessorR const &input, float *probs,
                                     int *indices, int batch_size, int length, int k, ffb_stream_t stream) {
    check(ffb_arg_topk(input.ptr, speculative_decoding ? probs : nullptr, indices, batch_size, length, k, input.data_type,
                       stream), "ffb_arg_topk");
  }
};

// Softmax::inference_kernel_wrapper (src/ops/kernels/softmax.cu:121-166)
struct Softmax {
  static void inference_kernel_wrapper(ffb_batch_config const *bc, GenericTensorAccessorR const &input,
                                       GenericTensorAccessorW const &output, int num_classes, ffb_stream_t stream) {
    check(ffb_softmax(input.ptr, output.ptr, bc->num_tokens, num_classes, input.data_type, stream), "ffb_softmax");
  }
};

// The D2H that ends a model step (ArgMax::inference_task, src/ops/argmax.cc:300-372)
inline void fetch_inference_result(int const *indices_dev, int num_tokens, ffb_inference_result *ir, ffb_stream_t stream) {
  check(ffb_fetch_inference_result(indices_dev, num_tokens, ir, stream), "ffb_fetch_inference_result");
}

// RequestManager (include/flexflow/request_manager.h:117-357): configuration, request registration and the batch
// builders, with the reference's method names; the serving loops are the callers' (see host/incr_decoding.cc).
class RequestManager {
public:
  void set_max_requests_per_batch(int n) { d_.max_requests_per_batch = n; }
  void set_max_tokens_per_batch(int n) { d_.max_tokens_per_batch = n; }
  void set_max_spec_tree_token_num(int n) { d_.max_spec_tree_token_num = n; }
  void set_max_sequence_length(int n) { d_.max_sequence_length = n; }
  int get_max_requests_per_batch() const { return d_.max_requests_per_batch; }
  int get_max_tokens_per_batch() const { return d_.max_tokens_per_batch; }
  int get_max_verify_tokens_per_batch() const {
    return d_.max_tokens_per_batch + d_.max_spec_tree_token_num * d_.max_requests_per_batch;
  }
  int register_ssm_model() { return d_.num_ssms++; }
  void push_spec_infer_tree_width(int w) { widths_.push_back(w); }
  void register_tokenizer(std::vector<int> const &eos_token_ids) {
    d_.num_eos_tokens = 0;
    for (int e : eos_token_ids) {
      if (d_.num_eos_tokens < 8) {
        d_.eos_token_ids[d_.num_eos_tokens++] = e;
      }
    }
  }
  ~RequestManager() { ffb_rm_destroy(h_); }
  RequestManager() = default;
  RequestManager(RequestManager const &) = delete;
  RequestManager &operator=(RequestManager const &) = delete;

  size_t register_new_request(std::vector<int> const &prompt_tokens, int max_length = -1, int max_new_tokens = -1) {
    size_t guid = 0;
    check(ffb_rm_register_request(handle(), prompt_tokens.data(), (int)prompt_tokens.size(), max_length, max_new_tokens, &guid),
          "ffb_rm_register_request");
    return guid;
  }
  void prepare_next_batch(ffb_batch_config const &old_bc, ffb_inference_result const &result, ffb_batch_config *new_bc) {
    check(ffb_rm_prepare_next_batch(handle(), &old_bc, &result, new_bc), "ffb_rm_prepare_next_batch");
  }
  void prepare_next_batch_init(ffb_tree_verify_batch_config const &old_bc, ffb_inference_result const &result, int model_id,
                               ffb_beam_search_batch_config *new_bc) {
    check(ffb_rm_prepare_next_batch_init(handle(), &old_bc, &result, model_id, new_bc), "ffb_rm_prepare_next_batch_init");
  }
  void prepare_next_batch_beam(ffb_beam_search_batch_config const &old_bc, ffb_beam_inference_result const &result,
                               ffb_beam_search_batch_config *new_bc) {
    check(ffb_rm_prepare_next_batch_beam(handle(), &old_bc, &result, new_bc), "ffb_rm_prepare_next_batch_beam");
  }
  void prepare_next_batch_verify(std::vector<ffb_beam_search_batch_config const *> const &old_batches,
                                 ffb_tree_verify_batch_config *new_bc) {
    check(ffb_rm_prepare_next_batch_verify(handle(), old_batches.data(), (int)old_batches.size(), new_bc),
          "ffb_rm_prepare_next_batch_verify");
  }
  std::vector<int> request_tokens(size_t guid) {
    int n = 0;
    check(ffb_rm_request_tokens(handle(), guid, nullptr, 0, &n), "ffb_rm_request_tokens");
    std::vector<int> t((size_t)n);
    check(ffb_rm_request_tokens(handle(), guid, t.data(), n, &n), "ffb_rm_request_tokens");
    return t;
  }
  int num_pending() { return ffb_rm_num_pending(handle()); }
  size_t get_num_processed_requests() { return ffb_rm_num_processed_requests(handle()); }

private:
  ffb_rm_t handle() {
    if (!h_) {
      check(ffb_rm_create(&d_, &h_), "ffb_rm_create");
      for (int w : widths_) {
        check(ffb_rm_push_spec_infer_tree_width(h_, w), "ffb_rm_push_spec_infer_tree_width");
      }
    }
    return h_;
  }
  ffb_rm_desc d_{};
  std::vector<int> widths_;
  ffb_rm_t h_ = nullptr;
};

// FileDataLoader (include/flexflow/utils/file_loader.h, src/runtime/file_loader.cc) for the layers of the path
class FileDataLoader {
public:
  FileDataLoader(std::string weights_folder, int num_heads, int num_kv_heads, int hidden_dim, int qkv_inner_dim,
                 int tensor_parallelism_degree, int data_type)
      : folder_(std::move(weights_folder)), nh_(num_heads), nkv_(num_kv_heads), hidden_(hidden_dim), dh_(qkv_inner_dim),
        tp_(tensor_parallelism_degree), dt_(data_type) {}
  // qkv_proj ([3 * H_l][hidden]) or o_proj ([hidden][H_l]) weight of one TP shard into host memory
  void load_attention_weights_to_dense(std::string const &layer_name, bool load_o_proj, int tp_rank, void *dst_host) const {
    check(ffb_load_attention_weights_to_dense(folder_.c_str(), layer_name.c_str(), dt_, nh_, nkv_, hidden_, dh_, tp_, tp_rank,
                                              load_o_proj, dst_host), "ffb_load_attention_weights_to_dense");
  }
  void load_from_file(std::string const &name, void *dst_host, size_t count) const {
    check(ffb_load_from_file((folder_ + "/" + name).c_str(), dst_host, count, dt_), "ffb_load_from_file");
  }
  void load_from_quantized_file(std::string const &name, int quantization_type, size_t blob_bytes, void *blob_host) const {
    check(ffb_load_from_quantized_file((folder_ + "/" + name).c_str(), quantization_type, dt_, blob_bytes, blob_host),
          "ffb_load_from_quantized_file");
  }

private:
  std::string folder_;
  int nh_, nkv_, hidden_, dh_, tp_, dt_;
};

} // namespace ffb
