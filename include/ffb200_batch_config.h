/* ffb200_batch_config.h — byte-compatible C mirrors of FlexFlow Serve's batch descriptors.
 *
 * These PODs are the drop-in ABI: a Legion task that holds a FlexFlow::BatchConfig (or
 * TreeVerifyBatchConfig / BeamSearchBatchConfig) can reinterpret_cast its pointer to the
 * matching struct below and hand it to the C entry points of ffb200.h.
 *
 * Reference layouts (all under /root/reference):
 *   include/flexflow/batch_config.h:53-164   BatchConfig, PerRequestInfo, PerTokenInfo, BitMask
 *   include/flexflow/batch_config.h:138-158  TreeVerifyBatchConfig::CommittedTokensInfo
 *   include/flexflow/batch_config.h:166-240  BeamSearchBatchConfig, InferenceResult
 *   include/flexflow/config.h:72-89          CombinedBatchConfigMetaStruct (device mirror)
 *   include/flexflow/ffconst.h:29-39,75-79   DataType / InferenceMode enum values
 *
 * The reference classes are polymorphic (one vptr) and the derived classes re-use the base's
 * tail padding (Itanium ABI), so the derived mirrors are written out flat here instead of
 * nesting the base. Sizes/offsets are pinned by the _Static_asserts at the bottom against the
 * numbers measured from the reference headers with nvcc 12.9 / gcc 13 on x86-64.
 */
#ifndef FFB200_BATCH_CONFIG_H
#define FFB200_BATCH_CONFIG_H

#include <stddef.h>
#include <stdint.h>
#include <stdbool.h>

#ifdef __cplusplus
extern "C" {
#endif

/* include/flexflow/batch_config.h:75-78 */
#define FFB_MAX_NUM_REQUESTS 65
#define FFB_MAX_NUM_TOKENS 1024
#define FFB_MAX_SPEC_TREE_TOKEN_NUM 64
#define FFB_MAX_PEFT_CONFIG_SIZE 1024
/* include/flexflow/batch_config.h:196-200 */
#define FFB_MAX_BEAM_WIDTH 3
#define FFB_MAX_BEAM_DEPTH 8
#define FFB_MAX_SPECULATIVE_TREE_BRANCHES 3
#define FFB_BEAM_TOKEN_INFO_LEN (FFB_MAX_NUM_TOKENS + FFB_MAX_SPEC_TREE_TOKEN_NUM * FFB_MAX_NUM_REQUESTS)

/* include/flexflow/ffconst.h:29-39 (DataType) */
enum { FFB_DT_HALF = 43, FFB_DT_FLOAT = 44, FFB_DT_INT4 = 46, FFB_DT_INT8 = 47, FFB_DT_NONE = 49,
       /* extension, not in the reference (it has no bf16) */
       FFB_DT_BF16 = 143 };
/* include/flexflow/ffconst.h:75-79 (InferenceMode) */
enum { FFB_INC_DECODING_MODE = 2001, FFB_BEAM_SEARCH_MODE = 2002, FFB_TREE_VERIFY_MODE = 2003 };

/* BatchConfig::PerRequestInfo, batch_config.h:86-115 — 1072 bytes */
typedef struct {
  int first_token_depth_in_request;
  int first_token_offset_in_batch;
  int num_tokens_in_batch;
  int max_length;
  int batch_config_request_id; /* compaction table: entry k holds the slot of the k-th active request */
  bool prompt_phase;
  size_t request_guid;
  size_t peft_model_id;
  char peft_model_config_str[FFB_MAX_PEFT_CONFIG_SIZE];
  bool peft_bwd;
  bool optimizer_tasks[4];
} ffb_per_request_info;

/* BatchConfig::PerTokenInfo, batch_config.h:116-120 — 12 bytes */
typedef struct {
  int abs_depth_in_request;
  int request_index;
  int token_id;
} ffb_per_token_info;

/* BatchConfig::BitMask, batch_config.h:122-137 — 528 bytes.
 * mask[k] bit q set  <=>  tree key k is visible to tree query q. */
typedef struct {
  unsigned long long mask[FFB_MAX_SPEC_TREE_TOKEN_NUM];
  int non_tree_cache_size;
  int tree_size;
  int this_layer_size;
  int prompt_size;
} ffb_bit_mask;

/* TreeVerifyBatchConfig::CommittedTokensInfo, batch_config.h:151-155 */
typedef struct {
  int token_index;   /* index of the token in the PREVIOUS verify batch */
  int request_index; /* request slot */
  int token_depth;   /* destination position in the request's sequence */
} ffb_committed_token_info;

/* BeamSearchBatchConfig::BeamSearchPerRequestInfo, batch_config.h:204-215 — 52 bytes */
typedef struct {
  int beam_size;
  int current_depth;
  int max_depth;
  int tokens[FFB_MAX_SPECULATIVE_TREE_BRANCHES];
  float probs[FFB_MAX_SPECULATIVE_TREE_BRANCHES];
  int parent_id[FFB_MAX_SPECULATIVE_TREE_BRANCHES];
  int sub_request_num;
} ffb_beam_per_request_info;

typedef struct {
  int sub_request_index;
} ffb_beam_per_token_info;

#define FFB_BATCH_CONFIG_FIELDS                                   \
  void *_vptr; /* the reference class is polymorphic */          \
  int num_tokens, num_peft_tokens, num_peft_label_tokens;         \
  int num_generation_tokens;                                      \
  ffb_bit_mask causalMask[FFB_MAX_NUM_REQUESTS];                  \
  ffb_per_request_info requestsInfo[FFB_MAX_NUM_REQUESTS];        \
  ffb_per_token_info tokensInfo[FFB_MAX_NUM_TOKENS];              \
  ffb_per_token_info labelsInfo[FFB_MAX_NUM_TOKENS];              \
  bool request_completed[FFB_MAX_NUM_REQUESTS];                   \
  bool request_running[FFB_MAX_NUM_REQUESTS];

/* BatchConfig — 128736 bytes */
typedef struct { FFB_BATCH_CONFIG_FIELDS } ffb_batch_config;

/* TreeVerifyBatchConfig — 141024 bytes (num_tokens_to_commit lives in the base's tail padding) */
typedef struct {
  FFB_BATCH_CONFIG_FIELDS
  int num_tokens_to_commit;
  ffb_committed_token_info committed_tokens[FFB_MAX_NUM_TOKENS];
} ffb_tree_verify_batch_config;

/* BeamSearchBatchConfig — 153144 bytes */
typedef struct {
  FFB_BATCH_CONFIG_FIELDS
  size_t beam_width;
  size_t target_iterations;
  int speculative_request_num;
  int model_id;
  ffb_beam_per_request_info beamRequestsInfo[FFB_MAX_NUM_REQUESTS];
  ffb_beam_per_token_info beamTokenInfo[FFB_BEAM_TOKEN_INFO_LEN];
  int sub_requests[FFB_MAX_NUM_REQUESTS];
  size_t current_iteration;
} ffb_beam_search_batch_config;

/* CombinedBatchConfigMetaStruct, config.h:72-89 — the device-resident mirror the op Metas alias
 * (handle.batch_config_metadata). 152760 bytes. */
typedef struct {
  ffb_per_token_info tokens_info[FFB_MAX_NUM_TOKENS];
  ffb_per_request_info requestsInfo[FFB_MAX_NUM_REQUESTS];
  ffb_bit_mask causalMask[FFB_MAX_NUM_REQUESTS];
  bool request_completed[FFB_MAX_NUM_REQUESTS];
  ffb_beam_per_token_info beamTokenInfo[FFB_BEAM_TOKEN_INFO_LEN];
  ffb_beam_per_request_info beamRequestsInfo[FFB_MAX_NUM_REQUESTS];
  ffb_committed_token_info committed_tokens[FFB_MAX_NUM_TOKENS];
} ffb_combined_batch_config_meta;

/* InferenceResult, batch_config.h:166-170 */
typedef struct {
  int token_ids[FFB_MAX_NUM_TOKENS];
  float finetuning_loss;
} ffb_inference_result;

#ifdef __cplusplus
}
#define FFB_SA(c, m) static_assert(c, m)
#else
#define FFB_SA(c, m) _Static_assert(c, m)
#endif

FFB_SA(sizeof(ffb_per_request_info) == 1072, "PerRequestInfo");
FFB_SA(sizeof(ffb_per_token_info) == 12, "PerTokenInfo");
FFB_SA(sizeof(ffb_bit_mask) == 528, "BitMask");
FFB_SA(sizeof(ffb_beam_per_request_info) == 52, "BeamSearchPerRequestInfo");
FFB_SA(sizeof(ffb_batch_config) == 128736, "BatchConfig");
FFB_SA(offsetof(ffb_batch_config, causalMask) == 24, "BatchConfig.causalMask");
FFB_SA(offsetof(ffb_batch_config, requestsInfo) == 34344, "BatchConfig.requestsInfo");
FFB_SA(offsetof(ffb_batch_config, tokensInfo) == 104024, "BatchConfig.tokensInfo");
FFB_SA(offsetof(ffb_batch_config, request_completed) == 128600, "BatchConfig.request_completed");
FFB_SA(sizeof(ffb_tree_verify_batch_config) == 141024, "TreeVerifyBatchConfig");
FFB_SA(offsetof(ffb_tree_verify_batch_config, num_tokens_to_commit) == 128732, "num_tokens_to_commit");
FFB_SA(offsetof(ffb_tree_verify_batch_config, committed_tokens) == 128736, "committed_tokens");
FFB_SA(sizeof(ffb_beam_search_batch_config) == 153144, "BeamSearchBatchConfig");
FFB_SA(offsetof(ffb_beam_search_batch_config, beamRequestsInfo) == 128760, "beamRequestsInfo");
FFB_SA(sizeof(ffb_combined_batch_config_meta) == 152760, "CombinedBatchConfigMetaStruct");

#endif /* FFB200_BATCH_CONFIG_H */
