#!/bin/sh
# lift_host.sh REF OUT_CLASS_INC OUT_FUNCS_INC — streams the reference's host-side serving-control text into two temporary
# include files for oracle/ref_host_wrap.cpp (see its header for the list and line ranges).  Test infrastructure; the
# temporaries are deleted by the Makefile after the compile.
set -e
REF="$1"; CLS="$2"; FN="$3"
HERE="$(dirname "$0")"
RMH="$REF/include/flexflow/request_manager.h"
RMC="$REF/src/runtime/request_manager.cc"
BCC="$REF/src/runtime/batch_config.cc"
BSC="$REF/src/runtime/beam_search_batch_config.cc"
TVC="$REF/src/runtime/tree_verify_batch_config.cc"

# struct Request, struct BeamTree, class RequestManager: from "struct Request {" up to the namespace's closing line
awk '/^struct Request \{/ {p=1} /^}; \/\/ namespace FlexFlow/ {p=0} p {print}' "$RMH" > "$CLS"
test -s "$CLS"

: > "$FN"
lift() { # file, pattern
  awk -v pat="$2" -f "$HERE/lift_fn.awk" "$1" > "$FN.one"
  test -s "$FN.one" || { echo "lift_host.sh: nothing matched '$2' in $1" >&2; exit 1; }
  cat "$FN.one" >> "$FN"; echo >> "$FN"
}
lift "$BCC" '^void set_optimizer_tasks\('
lift "$BCC" '^BatchConfig::BatchConfig\(\)'
for f in num_active_requests num_active_tokens num_active_infr_tokens num_active_peft_tokens max_requests_per_batch \
         max_tokens_per_batch max_verify_tokens_per_batch max_sequence_length max_spec_tree_token_num; do
  lift "$BCC" "^int BatchConfig::$f\\(\\)"
done
grep -E '^#define DEFAULT_(BEAM_WIDTH|TARGET_ITERATIONS) ' "$BSC" >> "$FN"
lift "$BSC" '^BeamSearchBatchConfig::BeamSearchBatchConfig\(\) '
lift "$TVC" '^TreeVerifyBatchConfig::TreeVerifyBatchConfig\(\)'
lift "$RMC" '^RequestManager::RequestManager\(\)'
for f in set_max_requests_per_batch get_max_requests_per_batch set_max_tokens_per_batch set_max_spec_tree_token_num \
         get_max_tokens_per_batch get_max_spec_tree_token_num get_max_verify_tokens_per_batch set_max_sequence_length \
         get_max_sequence_length push_spec_infer_tree_width is_eos_token check_inf_req_completion; do
  lift "$RMC" "^[a-z]+ RequestManager::$f\\("
done
lift "$RMC" '^BatchConfig RequestManager::prepare_next_batch\(BatchConfig const &old_bc,'
lift "$RMC" 'RequestManager::prepare_next_batch_init\(TreeVerifyBatchConfig const &old_bc,'
lift "$RMC" 'RequestManager::prepare_next_batch_beam\(BeamSearchBatchConfig const &old_bc,'
lift "$RMC" '^TreeVerifyBatchConfig RequestManager::prepare_next_batch_verify\($'
for f in store_beam_metadata update_beam_metadata initBitMask updateBitMask appendPendingRequest appendBitMask; do
  lift "$RMC" "^void RequestManager::$f\\("
done
lift "$RMC" '^    RequestManager::traverse_verify_tree\('
lift "$RMC" '^    RequestManager::traverse_beam_tree\('
lift "$RMC" '^    RequestManager::merge_dfs_trees\('
rm -f "$FN.one"
