// ref_attn_wrap.cu — compiles the REFERENCE's own decode attention kernel,
// compute_attention_kernel_generation_kernel (/root/reference/src/ops/inc_multihead_self_attention.cu:372-623), with the
// reference's own headers (flexflow/batch_config.h, flexflow/ops/kernels/inc_multihead_self_attention_utils.cuh, read
// where they lie) into oracle/_ref/libref_attn.so.  TEST INFRASTRUCTURE ONLY: the GPU tests compare ffb_inc_attn_forward
// with it on the same inputs and tools/bench_attn.py times it beside ours.  Nothing is copied into this repository:
// the Makefile streams the kernel's text (from its `__global__` line to its closing brace) into a temporary file that
// is included below and deleted after the compile.  The rest of that .cu needs Legion-typed Metas, accessors, cuBLAS
// and cuDNN handles and is not buildable here.
//
// Supplied here instead of the .cu's surroundings: the template header of the kernel (:372-377), WARP_SIZE (:30),
// QKV_WEIGHT_NUM (flexflow/ops/kernels/inc_multihead_self_attention_kernels.h:4) and the launch, which restates
// LAUNCH_ATTENTION_SCORE_KERNEL / compute_attention_kernel_generation (:880-927): grid (num_q_heads,
// num_generation_tokens), 128 threads, THREADS_PER_KEY 4, THREADS_PER_VALUE Dh * sizeof(DT) / 16, dynamic shared memory
// smem_size_in_bytes<DT>(Dh, max_sequence_length, THREADS_PER_VALUE, 128).
#include <cassert>
#include <cfloat>
#include <cstdint>
#include <cuComplex.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <math_constants.h>

#include "flexflow/batch_config.h"
#include "flexflow/ops/kernels/inc_multihead_self_attention_utils.cuh"

#define WARP_SIZE 32
#define QKV_WEIGHT_NUM 3

namespace FlexFlow {
namespace Kernels {
namespace IncMultiHeadAttention {

template <typename DT, int THREADS_PER_BLOCK, int Dh, int Dh_MAX, int THREADS_PER_KEY, int THREADS_PER_VALUE>
#include REF_GEN_KERNEL_INC

// apply_rotary_embedding_hf (:664-738): the text from its name line to its closing brace; `template <typename DT>
// __global__ void` (:664-665) supplied here.  CUDA_KERNEL_LOOP as in include/flexflow/utils/cuda_helper.h:64-66.
#define CUDA_KERNEL_LOOP(i, n) \
  for (Legion::coord_t i = blockIdx.x * blockDim.x + threadIdx.x; i < (n); i += blockDim.x * gridDim.x)
template <typename DT>
__global__ void
#include REF_ROPE_KERNEL_INC

// compute_spec_inc_attention_kernel_generation_kernel (src/ops/spec_inc_multihead_self_attention.cu:36-309), same
// template header; launch restated from LAUNCH_SPEC_INC_ATTENTION_SCORE_KERNEL (:391-442): grid (num_q_heads,
// speculative_request_num), shared memory sized for max_sequence_length + max_spec_tree_token_num.
template <typename DT, int THREADS_PER_BLOCK, int Dh, int Dh_MAX, int THREADS_PER_KEY, int THREADS_PER_VALUE>
#include REF_SPEC_KERNEL_INC

// compute_attention_kernel_fused_kernel (src/ops/tree_inc_multihead_self_attention.cu:35-333), same template header.
// The launch restates LAUNCH_TREE_VERIFY_ATTENTION_SCORE_KERNEL (:494-573) and smem_size_in_bytes_tree
// (inc_multihead_self_attention_utils.cuh:494-534; restated because the header's version calls
// BatchConfig::max_requests_per_batch(), which lives in the runtime).  That size follows the QUERY length only (SURVEY
// F7) while the kernel indexes its logits by key position, so the launcher refuses batches whose longest request
// would write past the allocation: -2.
template <typename DT, int THREADS_PER_BLOCK, int Dh, int Dh_MAX, int THREADS_PER_KEY, int THREADS_PER_VALUE>
#include REF_TREE_KERNEL_INC

// commit_tokens_kernel (tree_...cu:335-372) and update_tree_branch_kv_cache_fused (:433-478): the tree-mode KV
// placement; `template <typename DT>` supplied here, launches restated from commit_tokens (:374-398, with the
// reference's parallelism of hidden * KV_WEIGHT_NUM * tokens) and compute_attention_kernel_fused (:536-555).
template <typename DT>
#include REF_COMMIT_KERNEL_INC
template <typename DT>
#include REF_UPDATE_KERNEL_INC

template <int Dh>
static int launch_tree_half(void const *qkv, void const *kc, void const *vc, void *out, float scale, int max_seq,
                            int max_tokens, int hidden, void *request_infos, int num_q_heads, int num_requests,
                            void *causal_mask, void *request_completed, int max_query_length, int max_total_length,
                            cudaStream_t stream) {
  constexpr int TPV = threads_per_value_t<half, Dh>::value;
  size_t const qk_sz = div_up(max_query_length + 1, 4) * 16; // utils.cuh:515-518
  size_t const softmax_sz = 2 * qk_sz;
  size_t const q_size = Dh * sizeof(half);
  size_t const red_sz = (128 / TPV) * Dh * sizeof(float) / 2;
  size_t const total = softmax_sz + red_sz + q_size; // shared_mem[1], :531-533
  if ((size_t)max_total_length * sizeof(float) > total) {
    return -2;
  }
  dim3 grid(num_q_heads, num_requests);
  compute_attention_kernel_fused_kernel<half, 128, Dh, Dh, 4, TPV><<<grid, 128, total, stream>>>(
      (half const *)qkv, (half const *)kc, (half const *)vc, (half *)out, scale, max_seq, max_tokens, Dh, hidden,
      (BatchConfig::PerRequestInfo *)request_infos, num_q_heads, num_requests, (BatchConfig::BitMask *)causal_mask,
      (bool *)request_completed, (int)qk_sz);
  return (int)cudaGetLastError();
}

template <int Dh>
static int launch_spec_half(void const *qkv, void const *kc, void const *vc, void *out, float scale, int max_seq, int hidden,
                            void *request_infos, void *beam_request_infos, void *causal_mask, void *request_completed,
                            int num_q_heads, int num_spec_requests, cudaStream_t stream) {
  constexpr int TPV = threads_per_value_t<half, Dh>::value;
  size_t const smem = smem_size_in_bytes<half>(Dh, max_seq, TPV, 128);
  auto kern = compute_spec_inc_attention_kernel_generation_kernel<half, 128, Dh, Dh, 4, TPV>;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  dim3 grid(num_q_heads, num_spec_requests);
  kern<<<grid, 128, smem, stream>>>((half const *)qkv, (half const *)kc, (half const *)vc, (half *)out, scale, max_seq, Dh,
                                    hidden, (BatchConfig::PerRequestInfo *)request_infos,
                                    (BeamSearchBatchConfig::BeamSearchPerRequestInfo *)beam_request_infos,
                                    (BatchConfig::BitMask *)causal_mask, (bool *)request_completed);
  return (int)cudaGetLastError();
}

template <int Dh>
static int launch_half(void const *qkv, void const *kc, void const *vc, void *out, float scale, int max_seq, int hidden,
                       void *request_infos, int num_q_heads, int num_gen_tokens, cudaStream_t stream) {
  constexpr int TPV = threads_per_value_t<half, Dh>::value;
  size_t const smem = smem_size_in_bytes<half>(Dh, max_seq, TPV, 128);
  auto kern = compute_attention_kernel_generation_kernel<half, 128, Dh, Dh, 4, TPV>;
  if (smem > 48 * 1024) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  }
  dim3 grid(num_q_heads, num_gen_tokens);
  kern<<<grid, 128, smem, stream>>>((half const *)qkv, (half const *)kc, (half const *)vc, (half *)out, scale, max_seq,
                                    Dh, hidden, (BatchConfig::PerRequestInfo *)request_infos);
  return (int)cudaGetLastError();
}

} // namespace IncMultiHeadAttention
} // namespace Kernels
} // namespace FlexFlow

extern "C" {

// query: devQKVProjArray [tokens][3][hidden] (generation tokens first); key/value cache: [request][max_seq][hidden];
// request_infos: device copy of BatchConfig::requestsInfo.  Returns a cudaError_t.
int ref_inc_generation_half(void const *qkv, void const *key_cache, void const *value_cache, void *out, float scale,
                            int max_seq, int head_dim, int hidden, void *request_infos_dev, int num_q_heads,
                            int num_gen_tokens, void *stream) {
  using namespace FlexFlow::Kernels::IncMultiHeadAttention;
  if (head_dim == 128) {
    return launch_half<128>(qkv, key_cache, value_cache, out, scale, max_seq, hidden, request_infos_dev, num_q_heads,
                            num_gen_tokens, (cudaStream_t)stream);
  }
  if (head_dim == 64) {
    return launch_half<64>(qkv, key_cache, value_cache, out, scale, max_seq, hidden, request_infos_dev, num_q_heads,
                           num_gen_tokens, (cudaStream_t)stream);
  }
  return -1;
}
// max_seq = max_sequence_length + max_spec_tree_token_num (the cache's row count in beam mode); the four descriptor
// arrays are device copies of BeamSearchBatchConfig::{requestsInfo, beamRequestsInfo, causalMask, request_completed}.
int ref_spec_generation_half(void const *qkv, void const *key_cache, void const *value_cache, void *out, float scale,
                             int max_seq, int head_dim, int hidden, void *request_infos_dev, void *beam_request_infos_dev,
                             void *causal_mask_dev, void *request_completed_dev, int num_q_heads, int num_spec_requests,
                             void *stream) {
  using namespace FlexFlow::Kernels::IncMultiHeadAttention;
  if (head_dim == 128) {
    return launch_spec_half<128>(qkv, key_cache, value_cache, out, scale, max_seq, hidden, request_infos_dev,
                                 beam_request_infos_dev, causal_mask_dev, request_completed_dev, num_q_heads,
                                 num_spec_requests, (cudaStream_t)stream);
  }
  if (head_dim == 64) {
    return launch_spec_half<64>(qkv, key_cache, value_cache, out, scale, max_seq, hidden, request_infos_dev,
                                beam_request_infos_dev, causal_mask_dev, request_completed_dev, num_q_heads,
                                num_spec_requests, (cudaStream_t)stream);
  }
  return -1;
}
// max_seq = max_sequence_length + max_spec_tree_token_num; max_query_length / max_total_length over the active requests
// as smem_size_in_bytes_tree computes them (num_tokens_in_batch, first_token_depth_in_request + num_tokens_in_batch).
int ref_tree_verify_half(void const *qkv, void const *key_cache, void const *value_cache, void *out, float scale,
                         int max_seq, int max_tokens_per_batch, int head_dim, int hidden, void *request_infos_dev,
                         int num_q_heads, int num_active_requests, void *causal_mask_dev, void *request_completed_dev,
                         int max_query_length, int max_total_length, void *stream) {
  using namespace FlexFlow::Kernels::IncMultiHeadAttention;
  if (head_dim == 128) {
    return launch_tree_half<128>(qkv, key_cache, value_cache, out, scale, max_seq, max_tokens_per_batch, hidden,
                                 request_infos_dev, num_q_heads, num_active_requests, causal_mask_dev,
                                 request_completed_dev, max_query_length, max_total_length, (cudaStream_t)stream);
  }
  if (head_dim == 64) {
    return launch_tree_half<64>(qkv, key_cache, value_cache, out, scale, max_seq, max_tokens_per_batch, hidden,
                                request_infos_dev, num_q_heads, num_active_requests, causal_mask_dev,
                                request_completed_dev, max_query_length, max_total_length, (cudaStream_t)stream);
  }
  return -1;
}
// prev_qkv: devQKVProjArray of the PREVIOUS verify batch; committed_dev: device copy of committed_tokens[].
int ref_commit_tokens_half(void const *prev_qkv, void *key_cache, void *value_cache, void const *committed_dev, int proj,
                           int num_tokens_to_commit, int num_active_tokens_in_last_batch, int max_seq, int hidden,
                           void *stream) {
  using namespace FlexFlow;
  if (num_tokens_to_commit <= 0) {
    return 0;
  }
  int const parallelism = hidden * 2 * num_tokens_to_commit;
  int blocks = (parallelism + 1023) / 1024;
  blocks = blocks > 32768 ? 32768 : blocks;
  Kernels::IncMultiHeadAttention::commit_tokens_kernel<half><<<blocks, parallelism < 1024 ? parallelism : 1024, 0, (cudaStream_t)stream>>>(
      (half const *)prev_qkv, (half *)key_cache, (half *)value_cache,
      (TreeVerifyBatchConfig::CommittedTokensInfo const *)committed_dev, proj, proj, proj, num_tokens_to_commit,
      num_active_tokens_in_last_batch, max_seq, hidden);
  return (int)cudaGetLastError();
}
int ref_update_tree_branch_kv_cache_half(void const *qkv, void *key_cache, void *value_cache, void const *token_infos_dev,
                                         void *request_infos_dev, int proj, int num_new_tokens, int max_seq, int hidden,
                                         void *stream) {
  using namespace FlexFlow;
  int const parallelism = hidden * num_new_tokens;
  int blocks = (parallelism + 1023) / 1024;
  blocks = blocks > 32768 ? 32768 : blocks;
  Kernels::IncMultiHeadAttention::update_tree_branch_kv_cache_fused<half><<<blocks, parallelism < 1024 ? parallelism : 1024, 0, (cudaStream_t)stream>>>(
      (half const *)qkv, (half *)key_cache, (half *)value_cache, (TreeVerifyBatchConfig::PerTokenInfo const *)token_infos_dev,
      (BatchConfig::PerRequestInfo *)request_infos_dev, proj, proj, proj, num_new_tokens, max_seq, hidden);
  return (int)cudaGetLastError();
}
int ref_sizeof_committed_tokens_info(void) { return (int)sizeof(FlexFlow::TreeVerifyBatchConfig::CommittedTokensInfo); }
int ref_sizeof_beam_per_request_info(void) { return (int)sizeof(FlexFlow::BeamSearchBatchConfig::BeamSearchPerRequestInfo); }
int ref_sizeof_bit_mask(void) { return (int)sizeof(FlexFlow::BatchConfig::BitMask); }
// compute_qkv_kernel's RoPE step (:838-857): in place on devQKVProjArray [tokens][3][hidden]; complex_scratch holds
// tokens * hidden cuFloatComplex; token_infos_dev is a device copy of BatchConfig::tokensInfo.
int ref_apply_rotary_embedding_half(void *qkv, void *complex_scratch, void const *token_infos_dev, float rope_theta,
                                    int llama3, float factor, float low_freq_factor, float high_freq_factor,
                                    int original_max_position_embeddings, int proj_size, int num_tokens, int hidden,
                                    void *stream) {
  using namespace FlexFlow;
  int const parallelism = num_tokens * hidden;
  int blocks = (parallelism + 1023) / 1024;
  blocks = blocks > 32768 ? 32768 : blocks;
  size_t const q_array_size = (size_t)num_tokens * hidden; // qProjSize * num_tokens * num_q_heads
  Kernels::IncMultiHeadAttention::apply_rotary_embedding_hf<half>
      <<<blocks, parallelism < 1024 ? parallelism : 1024, 0, (cudaStream_t)stream>>>(
          (half *)qkv, (cuFloatComplex *)complex_scratch, (BatchConfig::PerTokenInfo const *)token_infos_dev, rope_theta,
          llama3 != 0, factor, low_freq_factor, high_freq_factor, original_max_position_embeddings, proj_size, proj_size,
          num_tokens, q_array_size, hidden);
  return (int)cudaGetLastError();
}
int ref_sizeof_per_token_info(void) { return (int)sizeof(FlexFlow::BatchConfig::PerTokenInfo); }
int ref_sizeof_per_request_info(void) { return (int)sizeof(FlexFlow::BatchConfig::PerRequestInfo); }

} // extern "C"
