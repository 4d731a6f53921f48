/* ffb200.h — C ABI of the B200-native FlexFlow Serve attention hot path.
 *
 * Every entry point returns an int status (FFB_OK == 0), takes the CUDA stream explicitly as an
 * opaque pointer (cudaStream_t), never allocates on the hot call and never aborts; the reference
 * equivalents assert()/exit() instead (include/flexflow/utils/cuda_helper.h:12-60).
 * All pointers named *_dev are device pointers; *_host are host pointers.
 *
 * Each function cites the reference interface it replaces (paths under /root/reference).
 * The reference-side binding a FlexFlow maintainer would add is shown in INTEGRATION.md.
 */
#ifndef FFB200_H
#define FFB200_H

#include "ffb200_batch_config.h"

#ifdef __cplusplus
extern "C" {
#endif

#define FFB_OK 0
#define FFB_ERR_INVALID 1     /* bad argument / unsupported shape */
#define FFB_ERR_CUDA 2        /* a CUDA runtime call failed; see ffb_last_error() */
#define FFB_ERR_UNSUPPORTED 3 /* feature outside the hot path (e.g. PEFT backward) */
#define FFB_ERR_NCCL 4
#define FFB_ERR_STATE 5       /* call order violated (e.g. commit without a previous verify batch) */

typedef void *ffb_stream_t; /* cudaStream_t */

/* Human-readable description of the last failure on the calling thread. */
char const *ffb_last_error(void);
/* Library/ABI version and build info ("sm_100a", nvcc version). */
char const *ffb_version(void);
/* Device sanity check: FFB_OK only on compute capability 10.x. There is no CPU fallback. */
int ffb_check_device(void);
/* The FFB_* environment overrides (kernel selection for tests / A-B runs, profiling hooks) are read once and cached; a
 * process that changes one after the library has seen it calls this to have them read again. */
void ffb_reload_env(void);

/* ------------------------------------------------------------------------------------------
 * Attention operator state.  Plays the role of IncMultiHeadSelfAttentionMeta /
 * TreeIncMultiHeadSelfAttentionMeta / SpecIncMultiHeadSelfAttentionMeta
 * (include/flexflow/ops/inc_multihead_self_attention.h:139-202, ctor
 * src/ops/inc_multihead_self_attention.cu:1621-1807): owns the KV cache, the previous verify
 * batch's K/V (for commit) and split-KV scratch.  Head counts are the LOCAL (per TP shard) ones,
 * as computed in src/ops/inc_multihead_self_attention.cc:516-519.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int mode;               /* FFB_INC_DECODING_MODE / FFB_BEAM_SEARCH_MODE / FFB_TREE_VERIFY_MODE */
  int dtype;              /* FFB_DT_HALF or FFB_DT_BF16 (FFB_DT_FLOAT: see DESIGN.md) */
  int num_q_heads;        /* local query heads */
  int num_kv_heads;       /* local distinct KV heads (q heads per kv head = num_q_heads/num_kv_heads) */
  int head_dim;           /* qProjSize == kProjSize == vProjSize: 64 or 128 */
  int global_num_q_heads; /* for ALiBi position bias only */
  /* RotaryEmbeddingMeta, include/flexflow/inference.h:46-66 */
  int apply_rotary_embedding;
  float rope_theta;
  int rope_llama3; /* rope_type == "llama3" */
  float rope_factor;
  float rope_low_freq_factor;
  float rope_high_freq_factor;
  int rope_original_max_position_embeddings;
  /* FFModel::inc_multihead_self_attention flags, include/flexflow/model.h:738-753 */
  int scaling_query;
  float scaling_factor;
  int qk_prod_scaling;
  int position_bias;
  /* BatchConfig::max_requests_per_batch() etc. (src/runtime/batch_config.cc) */
  int max_requests;
  int max_sequence_length;
  int max_spec_tree_tokens; /* 0 in incremental mode; cache rows = max_seq + this */
  int max_tokens_per_batch;
} ffb_attn_desc;

typedef struct ffb_attn_s *ffb_attn_t;

/* new IncMultiHeadSelfAttentionMeta(...) — allocates device memory once. */
int ffb_attn_create(ffb_attn_desc const *desc, ffb_attn_t *out);
/* ~IncMultiHeadSelfAttentionMeta, inc_multihead_self_attention.cu:1809-1813 */
int ffb_attn_destroy(ffb_attn_t h);
/* Bytes of device memory held by the handle (KV cache + scratch). */
size_t ffb_attn_device_bytes(ffb_attn_t h);

/* m->profiling of the reference (src/ops/inc_multihead_self_attention.cu:1510-1538 records a
 * cudaEvent pair around the whole wrapper and printf()s): here the pair brackets the dominant
 * attention kernel only, and the elapsed time of the last forward call is read back explicitly.
 * enable = 2 is a measurement mode: forward calls launch ONLY the dominant attention kernel (no
 * RoPE / KV-append kernel, no events, no programmatic dependent launch) on the state the last full
 * call left behind - same descriptors, same cache rows, same output - so that a caller can bracket
 * a run of such launches with its own events and divide (bench.py: roofline.kernel_ms). */
int ffb_attn_set_profiling(ffb_attn_t h, int enable);
int ffb_attn_last_kernel_ms(ffb_attn_t h, float *ms_out);

/* Replaces RequestManager::load_batch_config_task (src/runtime/request_manager.cu:83-165):
 * packs the parts of the host batch descriptor the kernels need into a pinned staging block and
 * issues ONE cudaMemcpyAsync into the device mirror (the reference issues 2-6 pageable copies).
 * `bc_host` points at a BatchConfig / TreeVerifyBatchConfig / BeamSearchBatchConfig according
 * to `mode`.  `staging_pinned_host` must hold sizeof(ffb_combined_batch_config_meta) bytes.
 * Lifetime: the copy is asynchronous.  The call remembers, per (staging block, device), an event recorded behind the
 * copy and waits for it before it writes the same block again, so a caller may reuse ONE staging block for every step
 * however far its host thread runs ahead of the stream; bc_host itself is free again when the call returns.  The
 * staging block must stay allocated until the stream has drained. */
int ffb_load_batch_config(void const *bc_host, int mode, void *staging_pinned_host,
                          ffb_combined_batch_config_meta *meta_dev, ffb_stream_t stream);

/* IncMultiHeadSelfAttention::inference_kernel_wrapper
 * (include/flexflow/ops/inc_multihead_self_attention.h:115-119,
 *  src/ops/inc_multihead_self_attention.cu:944-985,1501-1539).
 * qkv_dev: [num_tokens][3][num_q_heads*head_dim] = [Q | K | V] per token, K/V at q-head granularity
 * out_dev: [num_tokens][num_q_heads*head_dim]. */
int ffb_inc_attn_forward(ffb_attn_t h, ffb_batch_config const *bc_host,
                         ffb_combined_batch_config_meta const *meta_dev, int shard_id,
                         void const *qkv_dev, void *out_dev, ffb_stream_t stream);

/* TreeIncMultiHeadSelfAttention::inference_kernel_wrapper
 * (include/flexflow/ops/tree_inc_multihead_self_attention.h:104-108,
 *  src/ops/tree_inc_multihead_self_attention.cu:575-626): commit accepted tokens of the previous
 * verify batch, RoPE, append the tree tokens' K/V, tree-masked attention. */
int ffb_tree_attn_forward(ffb_attn_t h, ffb_tree_verify_batch_config const *bc_host,
                          ffb_combined_batch_config_meta const *meta_dev, int shard_id,
                          void const *qkv_dev, void *out_dev, ffb_stream_t stream);

/* SpecIncMultiHeadSelfAttention::inference_kernel_wrapper
 * (include/flexflow/ops/spec_inc_multihead_self_attention.h:101-106,
 *  src/ops/spec_inc_multihead_self_attention.cu:681-724). */
int ffb_spec_attn_forward(ffb_attn_t h, ffb_beam_search_batch_config const *bc_host,
                          ffb_combined_batch_config_meta const *meta_dev, int shard_id,
                          void const *qkv_dev, void *out_dev, ffb_stream_t stream);

/* Test/diagnostic access to the private KV cache: copies rows [pos0, pos0+n) of request slot
 * `request` into k_out_dev/v_out_dev as [n][num_kv_heads][head_dim] (reference cache order,
 * src/ops/inc_multihead_self_attention.cu:1693-1712), or writes them from that layout. */
int ffb_attn_read_kv(ffb_attn_t h, int request, int pos0, int n, void *k_out_dev, void *v_out_dev,
                     ffb_stream_t stream);
int ffb_attn_write_kv(ffb_attn_t h, int request, int pos0, int n, void const *k_in_dev,
                      void const *v_in_dev, ffb_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Weight decompression.  Kernels::decompress_int4_general_weights / decompress_int8_general_weights
 * (src/ops/kernels/decompress_kernels.cu:27-70; launch src/ops/kernels/linear_kernels.cu:469-490).
 * blob layout: [values | offsets | scales]; element idx = o*in_dim + i; group = (o/32)*in_dim + i;
 * int4: byte idx/2, even idx = high nibble; w = DT(q) / scale + offset in DT arithmetic.
 * out_dev: [out_dim][in_dim] DT.  dtype: FFB_DT_HALF (bit-exact with the reference) or FFB_DT_BF16.
 * ------------------------------------------------------------------------------------------ */
size_t ffb_quant_blob_bytes(int quant_type, int dtype, int in_dim, int out_dim);
int ffb_dequant_int4(void const *blob_dev, void *out_dev, int in_dim, int out_dim, int dtype,
                     ffb_stream_t stream);
int ffb_dequant_int8(void const *blob_dev, void *out_dev, int in_dim, int out_dim, int dtype,
                     ffb_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Linear forward (QKV / o_proj dense layers).  Kernels::Linear::forward_kernel_wrapper /
 * inference_kernel_wrapper (src/ops/kernels/linear_kernels.cu:133-140,187-195,450-582):
 * out[batch][out_dim] = in[batch][in_dim] * W^T (+ bias), W stored [out_dim][in_dim] (in fastest).
 * quant_type FFB_DT_NONE: weight_dev is DT; FFB_DT_INT8 / FFB_DT_INT4: weight_dev is the blob
 * above.  Up to 64 tokens it is dequantised on chip inside the GEMM (in registers) and never materialised in HBM; above
 * that it is dequantised into a private scratch of in_dim x out_dim DT elements (allocated at create when max_batch >
 * 64) and fed to the tensor-core GEMM, as the reference does (linear_kernels.cu:464-528).  FFB_TCQ=1 selects an
 * in-shared-memory dequantising GEMM for those batches instead (measured slower; profiles/README.md).
 * Accumulation is fp32 (the reference accumulates in DT, linear_kernels.cu:507-528).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  int dtype;      /* FFB_DT_HALF / FFB_DT_BF16 */
  int quant_type; /* FFB_DT_NONE / FFB_DT_INT8 / FFB_DT_INT4 */
  int in_dim;
  int out_dim;
  int max_batch;  /* BatchConfig::max_tokens_per_batch() */
  int activation; /* 0 = AC_MODE_NONE (only mode on this path) */
} ffb_linear_desc;
typedef struct ffb_linear_s *ffb_linear_t;
int ffb_linear_create(ffb_linear_desc const *desc, ffb_linear_t *out);
int ffb_linear_destroy(ffb_linear_t h);
int ffb_linear_forward(ffb_linear_t h, void const *in_dev, void *out_dev, void const *weight_dev,
                       void const *bias_dev, int batch, ffb_stream_t stream);

/* ffb_linear_forward_static: no counterpart in the reference - added for
 * callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * out[batch][out_dim] = in[batch][in_dim] x weight[out_dim][in_dim]^T (+ bias), dense half / bf16 weights, computed
 * by the SM-pair tcgen05 GEMM with its split-K workspace in a file-scope __device__ array: the call needs no handle
 * and allocates nothing (no cudaMalloc / cudaFree, host or device side).  Limits: batch <= 1024, in_dim <= 16384 and
 * % 64 == 0, out_dim <= 16384 and % 16 == 0; anything larger is rejected with FFB_ERR_INVALID.  The scratch
 * (ffb_linear_static_scratch_bytes() = 19.4 MB) is held by every GPU that loads the library until exit and is SHARED
 * by all calls: calls whose kernels overlap on the GPU (different streams) race; calls on one stream are safe.
 * Results agree with ffb_linear_forward's within fp32-accumulation rounding (bit-identical where that entry point
 * dispatches to the same kernel and schedule, e.g. 512 x 4096 x 4096). */
int ffb_linear_forward_static(void const *in_dev, void *out_dev, void const *weight_dev, void const *bias_dev,
                              int batch, int in_dim, int out_dim, int dtype, ffb_stream_t stream);
size_t ffb_linear_static_scratch_bytes(void);

/* ------------------------------------------------------------------------------------------
 * AllReduce parallel op.  Kernels::AllReduce::inference_kernel_wrapper
 * (src/parallel_ops/kernels/allreduce_kernels.cu:56-78): sum over TP shards of
 * num_active_tokens*hidden elements.  Communicator setup replaces
 * src/runtime/model.cc:584-628 (ncclGetUniqueId task + ncclCommInitRank index task).
 * ------------------------------------------------------------------------------------------ */
#define FFB_NCCL_UNIQUE_ID_BYTES 128
typedef struct ffb_comm_s *ffb_comm_t;
int ffb_comm_get_unique_id(void *id_out_128B);
int ffb_comm_init_rank(void const *id_128B, int nranks, int rank, ffb_comm_t *out);
int ffb_comm_destroy(ffb_comm_t c);
/* Adopt an ncclComm_t the runtime already owns (FFHandler.ncclComm, include/flexflow/config.h:110). */
int ffb_comm_from_nccl(void *nccl_comm, int nranks, int rank, ffb_comm_t *out);
int ffb_allreduce(ffb_comm_t c, void const *in_dev, void *out_dev, size_t count, int dtype,
                  ffb_stream_t stream);
/* 1 when the communicator mapped every rank's receive area over NVLink (CUDA IPC) at creation. */
int ffb_comm_has_peer(ffb_comm_t c);
/* Linear (row-parallel o_proj) immediately followed by the AllReduce over the TP shards, as the reference runs them
 * back to back inside FusedOp::inference_task (src/ops/fused.cu:448-495; insertion rule src/runtime/model.cc:3421-3445).
 * Decode batches (<= 16 tokens) on a peer-mapped communicator run as ONE kernel that pushes the fp32 partial sums into
 * every rank's receive area and reduces them in rank order; otherwise ffb_linear_forward + ffb_allreduce.  Every
 * rank must make the same call.  A bias is added on every shard before the sum, as in the reference (each shard's
 * Linear adds it, linear_kernels.cu:531, then the AllReduce sums): the result carries nranks x bias on both paths. */
int ffb_linear_forward_allreduce(ffb_linear_t h, ffb_comm_t comm, void const *in_dev, void *out_dev,
                                 void const *weight_dev, void const *bias_dev, int batch, ffb_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * The two row operators either side of the chain that the reference's FusedOp spans.
 *
 * ffb_residual_rms_norm: Kernels::ResidualRMSNorm::inference_kernel_wrapper
 * (src/ops/kernels/residual_rms_norm_kernels.cu:97-164,205-232): residual_output = input1 + input2 (in DT),
 * rms = DT(rsqrt(mean(residual_output^2) + eps)), output = DT(DT(residual_output * rms) * weight).
 * input2_dev == NULL gives Kernels::RMSNorm::inference_kernel_wrapper (src/ops/kernels/rms_norm_kernels.cu:97-146)
 * on input1 (residual_output_dev is then not written).  Roundings and the fp32 summation order are the
 * reference's.  All tensors [num_tokens][dim] DT, weight [dim] DT; dim % 32 == 0, dim <= 16384.
 *
 * ffb_argmax: ArgMax::forward_kernel (src/ops/argmax.cu:64-102): indices[t] = first index of the maximum of
 * input[t][0..length); prob_dev (beam search: the maximum as float) and parent_dev (zeroed) may be NULL.
 * ------------------------------------------------------------------------------------------ */
int ffb_residual_rms_norm(void const *input1_dev, void const *input2_dev, void const *weight_dev,
                          void *residual_output_dev, void *output_dev, int num_tokens, int dim, float eps, int dtype,
                          ffb_stream_t stream);
int ffb_argmax(void const *input_dev, int *indices_dev, float *prob_dev, int *parent_dev, int num_tokens, int length,
               int dtype, ffb_stream_t stream);
/* ArgTopK::forward_kernel (src/ops/arg_topk.cu:373-449; the SSM's token selection in beam mode,
 * inference/models/llama.cc:277-295): indices[t][0..k) = the k largest entries of input[t][0..length), best first,
 * equal values in increasing index order; values_dev (speculative decoding: the entries as float) may be NULL.
 * The reference leaves values[t][k-1] unwritten (:330-331); here all k are written.  1 <= k <= 16. */
int ffb_arg_topk(void const *input_dev, float *values_dev, int *indices_dev, int num_tokens, int length, int k,
                 int dtype, ffb_stream_t stream);

/* Softmax::inference_kernel_wrapper (src/ops/kernels/softmax.cu:121-166,262-288; cudnnSoftmaxForward, ACCURATE, over the
 * class dimension): output[t][c] = exp(input[t][c] - max_t) / sum_c, fp32 inside, one rounding to DT.  The op between
 * lm_head and ArgMax / ArgTopK (inference/models/llama.cc:277-291).  In place (output_dev == input_dev) is allowed. */
int ffb_softmax(void const *input_dev, void *output_dev, int num_tokens, int num_classes, int dtype, ffb_stream_t stream);

/* ---- host-side SpecInfer tree logic (integer only; runs on the host, no CUDA call) ---------------------------------
 * The bit masks these build are what ffb_tree_attn_forward / ffb_spec_attn_forward read from causalMask[], and
 * ffb_traverse_verify_tree turns the LLM's per-node outputs into the accepted tokens and the commit list of the next
 * verify batch.  Replaces RequestManager::{initBitMask, updateBitMask, appendPendingRequest, appendBitMask,
 * traverse_verify_tree} (src/runtime/request_manager.cc:2382-2396, 2399-2427, 2430-2446, 2448-2517, 2583-2741).
 * Bits are set with 64-bit shifts (the reference's `1 << j` is a 32-bit shift, identical up to tree index 30). */
void ffb_bitmask_init(ffb_bit_mask *bm, int init_length);
void ffb_bitmask_update(ffb_bit_mask *bm, int init_length, int non_tree_size);
void ffb_bitmask_append_pending(ffb_bit_mask *bm, int init_length);
/* new_nodes = nodes of the new (deepest) layer; layer_nodes[i] = treeLayers[i].nodes_num_this_layer for
 * i < current_depth.  FFB_ERR_INVALID when the tree would exceed FFB_MAX_SPEC_TREE_TOKEN_NUM nodes (the reference
 * asserts), FFB_ERR_STATE when the layer widths do not add up to the nodes already in the mask. */
int ffb_bitmask_append(ffb_bit_mask *bm, int new_nodes, int const *layer_nodes, int current_depth);

typedef struct {
  int token_id;
  int depth; /* absolute position in the request's sequence */
} ffb_tree_node;
typedef struct {
  int depth;
  int index_in_batch;
} ffb_commit_entry;
/* input_tree: the speculated tree serialised in layer order (dfs_tree_inputs); llm_outputs[i]: the LLM's token for node
 * i with depth + 1; committed_in[i]: (depth, index in the verify batch) of node i.  Writes the verified (token, depth)
 * sequence and the commit list (at most n_in entries each).  Accept rule: node 0 always; node i > 0 iff it equals the
 * last verified (token, depth) and is the first such node or sits in the same slot of its layer as the first one.
 * FFB_ERR_INVALID if n_in < n_out, FFB_ERR_STATE if a committed depth disagrees with the tree (reference asserts). */
int ffb_traverse_verify_tree(ffb_tree_node const *input_tree, int n_in, ffb_tree_node const *llm_outputs, int n_out,
                             ffb_commit_entry const *committed_in, ffb_tree_node *verified, int *n_verified,
                             ffb_commit_entry *committed_out, int *n_committed);

/* ---- serving control: the batch builders of RequestManager -------------------------------------------
 * Host code, integer only.  One handle plays the role of the RequestManager singleton: the request table
 * (all_requests, pending_infr_request_queue), each request's speculation tree (Request::beam_trees), the serialised
 * tree handed to the verifier (dfs_tree_inputs) and the commit list (committed_tokens).  The four builders consume the
 * previous batch descriptor plus the model's result and fill the next descriptor in the reference's layout
 * (include/ffb200_batch_config.h), field for field what the reference's functions emit:
 *   ffb_rm_prepare_next_batch         RequestManager::prepare_next_batch        src/runtime/request_manager.cc:713-1138
 *   ffb_rm_prepare_next_batch_init    ...::prepare_next_batch_init              :1170-1580
 *   ffb_rm_prepare_next_batch_beam    ...::prepare_next_batch_beam (+ store_beam_metadata / update_beam_metadata
 *                                     :2217-2379)                                :1610-1892
 *   ffb_rm_prepare_next_batch_verify  ...::prepare_next_batch_verify (+ traverse_beam_tree / merge_dfs_trees
 *                                     :2743-2888)                                :1923-2215
 * One SpecInfer round is serve_spec_infer's (:3083-3173): init -> FFB_MAX_BEAM_DEPTH x (SSM step, beam) -> verify -> LLM
 * step.  Fields the reference leaves uninitialised (request_running in incremental mode, beam tokens / probs /
 * parent_id of slots that do not use them, ...) are zero here.  PEFT / fine-tuning requests are out of scope
 * (FFB_ERR_UNSUPPORTED never arises: they cannot be registered).  Where the reference asserts, these return
 * FFB_ERR_STATE / FFB_ERR_INVALID and leave *new_bc unspecified. */
typedef struct {
  int max_requests_per_batch;  /* RequestManager::set_max_requests_per_batch, request_manager.cc:117-122 */
  int max_tokens_per_batch;
  int max_spec_tree_token_num;
  int max_sequence_length;
  int num_ssms;                /* register_ssm_model calls (0: incremental decoding only) */
  int num_eos_tokens;          /* eos_token_ids, request_manager.cc:181-244 */
  int eos_token_ids[8];
} ffb_rm_desc;
typedef struct ffb_rm_s *ffb_rm_t;
enum { FFB_REQ_PENDING = 101, FFB_REQ_RUNNING = 102, FFB_REQ_COMPLETED = 103, FFB_REQ_FINISHING = 104 }; /* Request::Status */

/* BeamInferenceResult, include/flexflow/batch_config.h:231-240 */
typedef struct {
  int token_ids[FFB_MAX_NUM_TOKENS * FFB_MAX_SPECULATIVE_TREE_BRANCHES];
  float probs[FFB_MAX_NUM_TOKENS * FFB_MAX_SPECULATIVE_TREE_BRANCHES];
  int parent_id[FFB_MAX_NUM_TOKENS * FFB_MAX_SPECULATIVE_TREE_BRANCHES];
} ffb_beam_inference_result;

int ffb_rm_create(ffb_rm_desc const *desc, ffb_rm_t *out);
int ffb_rm_destroy(ffb_rm_t rm);
/* RequestManager::push_spec_infer_tree_width (request_manager.cc:168-171): beam width of SSM step k of every round. */
int ffb_rm_push_spec_infer_tree_width(ffb_rm_t rm, int tree_width);
/* RequestManager::register_new_request (request_manager.cc:334-441) for already tokenised prompts: max_length /
 * max_new_tokens follow its rules (-1 = unset; both unset: max_sequence_length - 1; max_new_tokens wins).  A prompt or
 * max_length that does not fit max_sequence_length gives FFB_ERR_INVALID (the reference returns INVALID_GUID). */
int ffb_rm_register_request(ffb_rm_t rm, int const *prompt_tokens, int num_tokens, int max_length, int max_new_tokens,
                            size_t *guid_out);
int ffb_rm_prepare_next_batch(ffb_rm_t rm, ffb_batch_config const *old_bc, ffb_inference_result const *result,
                              ffb_batch_config *new_bc);
int ffb_rm_prepare_next_batch_init(ffb_rm_t rm, ffb_tree_verify_batch_config const *old_bc,
                                   ffb_inference_result const *result, int model_id,
                                   ffb_beam_search_batch_config *new_bc);
int ffb_rm_prepare_next_batch_beam(ffb_rm_t rm, ffb_beam_search_batch_config const *old_bc,
                                   ffb_beam_inference_result const *result, ffb_beam_search_batch_config *new_bc);
int ffb_rm_prepare_next_batch_verify(ffb_rm_t rm, ffb_beam_search_batch_config const *const *old_batches,
                                     int num_batches, ffb_tree_verify_batch_config *new_bc);
/* Request state: status (Request::Status), the token sequence so far (copies min(n, capacity) ids, *num_tokens = n),
 * profiling counters (ProfileInfo::llm_decoding_steps / ssm_decoding_steps). */
int ffb_rm_request_status(ffb_rm_t rm, size_t guid, int *status_out);
int ffb_rm_request_tokens(ffb_rm_t rm, size_t guid, int *tokens_out, int capacity, int *num_tokens);
int ffb_rm_request_steps(ffb_rm_t rm, size_t guid, int *llm_decoding_steps, int *ssm_decoding_steps);
int ffb_rm_num_pending(ffb_rm_t rm);
size_t ffb_rm_num_processed_requests(ffb_rm_t rm);

/* The D2H that ends a model step: ArgMax::inference_task / ArgTopK::inference_speculative_task copy the
 * token ids (and, in beam mode, the probabilities) of the step into the InferenceResult the next builder reads
 * (src/ops/argmax.cc:300-372, src/ops/arg_topk.cc:395-437).  One (two) cudaMemcpyAsync on `stream` followed by a
 * stream synchronise - the call returns with the host struct filled, like the reference's copy_tensor_dev_to_host.
 * parent_id, which the reference leaves unwritten, is zero-filled for the entries copied. */
int ffb_fetch_inference_result(int const *indices_dev, int num_tokens, ffb_inference_result *result_host,
                               ffb_stream_t stream);
int ffb_fetch_beam_inference_result(int const *indices_dev, float const *probs_dev, int num_tokens, int k,
                                    ffb_beam_inference_result *result_host, ffb_stream_t stream);

/* ---- weight file formats and the quantiser (host code) ----------------------------------------------
 * ffb_load_from_file: load_from_file (src/runtime/file_loader.cc:345-377): a flat binary of `count` elements of
 * `dtype` (FFB_DT_HALF / FFB_DT_BF16 / FFB_DT_FLOAT); a short or missing file is FFB_ERR_INVALID (the reference
 * asserts).
 * ffb_load_attention_weights_to_dense: load_attention_weights_to_dense_v2 (file_loader.cc:215-343) for the dense layers
 * either side of the attention operator.  load_o_proj == 0: reads `<layer_name>.q_proj.weight`, `.k_proj.weight`,
 * `.v_proj.weight` ([heads * head_dim][hidden], HuggingFace order) and writes the column-parallel qkv_proj weight of
 * shard `tp_rank`: [3 * (num_heads / tp) * head_dim][hidden] = [Q_r | K_r | V_r] with the kv heads replicated to q-head
 * granularity (the layout ffb_*_attn_forward's input comes in).  load_o_proj != 0: reads `<layer_name>.o_proj.weight`
 * ([hidden][num_heads * head_dim]) and writes the row-parallel shard [hidden][(num_heads / tp) * head_dim].
 * tp_rank == -1 writes what the reference's function writes: all shards back to back (qkv) / the file as stored (o). */
int ffb_load_from_file(char const *path, void *dst_host, size_t count, int dtype);
int ffb_load_attention_weights_to_dense(char const *weights_folder, char const *layer_name, int dtype, int num_heads,
                                        int num_kv_heads, int hidden_dim, int head_dim, int tp_degree, int tp_rank,
                                        int load_o_proj, void *dst_host);
/* The quantiser of inference/utils/compress_llama_weights.py:15-58 (num_bits 8 / 4, group_size 32, group_dim 0,
 * asymmetric) on a [out_dim][in_dim] fp32 matrix: per (32-row group, column) min and scale = B / (max - min) in fp32,
 * q = round_half_even(clamp((w - min) * scale, 0, B)); writes the blob [values | offsets | scales] (ffb_quant_blob_bytes
 * bytes) with the group meta rounded to `dtype` (FFB_DT_HALF / FFB_DT_BF16).  A constant group gets q = 0 and an
 * infinite scale, which dequantises to the constant (the script would produce NaN casts there). */
int ffb_quantize_weights(float const *w_host, int in_dim, int out_dim, int quant_type, int dtype, void *blob_host);
/* The three files FileDataLoader::load_quantization_weight reads for a non-attention tensor (file_loader.cc:552-669):
 * `path` (ONE byte per element, 4-bit values included), `path`_offset, `path`_scale.  ffb_load_from_quantized_file is
 * load_from_quantized_file: it fills the blob, packing two 4-bit values per byte, the first in the high nibble. */
int ffb_save_quantized_file(char const *path, void const *blob_host, int in_dim, int out_dim, int quant_type, int dtype);
int ffb_load_from_quantized_file(char const *path, int quant_type, int dtype, size_t blob_bytes, void *blob_host);

#ifdef __cplusplus
}
#endif
#endif /* FFB200_H */
