"""ctypes mirrors of include/ffb200_batch_config.h (== FlexFlow's BatchConfig family).

Reference: include/flexflow/batch_config.h:53-240, include/flexflow/config.h:72-89.
Sizes are asserted at import against the numbers pinned in the C header.
"""
import ctypes as C

MAX_NUM_REQUESTS = 65
MAX_NUM_TOKENS = 1024
MAX_SPEC_TREE_TOKEN_NUM = 64
MAX_PEFT_CONFIG_SIZE = 1024
MAX_BEAM_WIDTH = 3
MAX_BEAM_DEPTH = 8
MAX_SPECULATIVE_TREE_BRANCHES = 3
BEAM_TOKEN_INFO_LEN = MAX_NUM_TOKENS + MAX_SPEC_TREE_TOKEN_NUM * MAX_NUM_REQUESTS

DT_HALF, DT_FLOAT, DT_INT4, DT_INT8, DT_NONE, DT_BF16 = 43, 44, 46, 47, 49, 143
INC_DECODING_MODE, BEAM_SEARCH_MODE, TREE_VERIFY_MODE = 2001, 2002, 2003


class PerRequestInfo(C.Structure):
    _fields_ = [
        ("first_token_depth_in_request", C.c_int),
        ("first_token_offset_in_batch", C.c_int),
        ("num_tokens_in_batch", C.c_int),
        ("max_length", C.c_int),
        ("batch_config_request_id", C.c_int),
        ("prompt_phase", C.c_bool),
        ("request_guid", C.c_size_t),
        ("peft_model_id", C.c_size_t),
        ("peft_model_config_str", C.c_char * MAX_PEFT_CONFIG_SIZE),
        ("peft_bwd", C.c_bool),
        ("optimizer_tasks", C.c_bool * 4),
    ]


class PerTokenInfo(C.Structure):
    _fields_ = [
        ("abs_depth_in_request", C.c_int),
        ("request_index", C.c_int),
        ("token_id", C.c_int),
    ]


class BitMask(C.Structure):
    _fields_ = [
        ("mask", C.c_ulonglong * MAX_SPEC_TREE_TOKEN_NUM),
        ("non_tree_cache_size", C.c_int),
        ("tree_size", C.c_int),
        ("this_layer_size", C.c_int),
        ("prompt_size", C.c_int),
    ]


class CommittedTokensInfo(C.Structure):
    _fields_ = [("token_index", C.c_int), ("request_index", C.c_int), ("token_depth", C.c_int)]


class BeamSearchPerRequestInfo(C.Structure):
    _fields_ = [
        ("beam_size", C.c_int),
        ("current_depth", C.c_int),
        ("max_depth", C.c_int),
        ("tokens", C.c_int * MAX_SPECULATIVE_TREE_BRANCHES),
        ("probs", C.c_float * MAX_SPECULATIVE_TREE_BRANCHES),
        ("parent_id", C.c_int * MAX_SPECULATIVE_TREE_BRANCHES),
        ("sub_request_num", C.c_int),
    ]


class BeamSearchPerTokenInfo(C.Structure):
    _fields_ = [("sub_request_index", C.c_int)]


_BASE_FIELDS = [
    ("_vptr", C.c_void_p),
    ("num_tokens", C.c_int),
    ("num_peft_tokens", C.c_int),
    ("num_peft_label_tokens", C.c_int),
    ("num_generation_tokens", C.c_int),
    ("causalMask", BitMask * MAX_NUM_REQUESTS),
    ("requestsInfo", PerRequestInfo * MAX_NUM_REQUESTS),
    ("tokensInfo", PerTokenInfo * MAX_NUM_TOKENS),
    ("labelsInfo", PerTokenInfo * MAX_NUM_TOKENS),
    ("request_completed", C.c_bool * MAX_NUM_REQUESTS),
    ("request_running", C.c_bool * MAX_NUM_REQUESTS),
]


class _BCMixin:
    """Host helpers restating src/runtime/batch_config.cc accessors."""

    def reset(self):
        C.memset(C.addressof(self), 0, C.sizeof(self))
        for i in range(MAX_NUM_REQUESTS):
            self.request_completed[i] = True
            self.requestsInfo[i].batch_config_request_id = -1
        return self

    def num_active_requests(self):
        return sum(0 if self.request_completed[i] else 1 for i in range(MAX_NUM_REQUESTS))

    def num_active_tokens(self):
        return self.num_tokens


class BatchConfig(C.Structure, _BCMixin):
    _fields_ = list(_BASE_FIELDS)


class TreeVerifyBatchConfig(C.Structure, _BCMixin):
    _fields_ = list(_BASE_FIELDS) + [
        ("num_tokens_to_commit", C.c_int),
        ("committed_tokens", CommittedTokensInfo * MAX_NUM_TOKENS),
    ]


class BeamSearchBatchConfig(C.Structure, _BCMixin):
    _fields_ = list(_BASE_FIELDS) + [
        ("beam_width", C.c_size_t),
        ("target_iterations", C.c_size_t),
        ("speculative_request_num", C.c_int),
        ("model_id", C.c_int),
        ("beamRequestsInfo", BeamSearchPerRequestInfo * MAX_NUM_REQUESTS),
        ("beamTokenInfo", BeamSearchPerTokenInfo * BEAM_TOKEN_INFO_LEN),
        ("sub_requests", C.c_int * MAX_NUM_REQUESTS),
        ("current_iteration", C.c_size_t),
    ]


class CombinedBatchConfigMeta(C.Structure):
    _fields_ = [
        ("tokens_info", PerTokenInfo * MAX_NUM_TOKENS),
        ("requestsInfo", PerRequestInfo * MAX_NUM_REQUESTS),
        ("causalMask", BitMask * MAX_NUM_REQUESTS),
        ("request_completed", C.c_bool * MAX_NUM_REQUESTS),
        ("beamTokenInfo", BeamSearchPerTokenInfo * BEAM_TOKEN_INFO_LEN),
        ("beamRequestsInfo", BeamSearchPerRequestInfo * MAX_NUM_REQUESTS),
        ("committed_tokens", CommittedTokensInfo * MAX_NUM_TOKENS),
    ]


class InferenceResult(C.Structure):
    """include/flexflow/batch_config.h:166-170"""
    _fields_ = [("token_ids", C.c_int * MAX_NUM_TOKENS), ("finetuning_loss", C.c_float)]


class BeamInferenceResult(C.Structure):
    """include/flexflow/batch_config.h:231-240"""
    _fields_ = [("token_ids", C.c_int * (MAX_NUM_TOKENS * MAX_SPECULATIVE_TREE_BRANCHES)),
                ("probs", C.c_float * (MAX_NUM_TOKENS * MAX_SPECULATIVE_TREE_BRANCHES)),
                ("parent_id", C.c_int * (MAX_NUM_TOKENS * MAX_SPECULATIVE_TREE_BRANCHES))]


assert C.sizeof(PerRequestInfo) == 1072
assert C.sizeof(BitMask) == 528
assert C.sizeof(BatchConfig) == 128736
assert C.sizeof(TreeVerifyBatchConfig) == 141024
assert TreeVerifyBatchConfig.num_tokens_to_commit.offset == 128732
assert C.sizeof(BeamSearchBatchConfig) == 153144
assert C.sizeof(CombinedBatchConfigMeta) == 152760
