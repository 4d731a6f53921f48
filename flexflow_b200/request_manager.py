"""Host mirror of the reference's RequestManager (include/flexflow/request_manager.h:117-357) over the C ABI's ffb_rm_*
entry points (include/ffb200.h): same method names and argument meaning; the batch descriptors are the ctypes mirrors of
flexflow_b200.batch_config.  A failed reference assert surfaces as lib.FFBError."""
import ctypes as C

from . import batch_config as B
from . import lib as _lib

PENDING, RUNNING, COMPLETED, FINISHING = 101, 102, 103, 104  # Request::Status


class RequestManager:
    def __init__(self):
        self._h = None
        self.max_requests_per_batch = self.max_tokens_per_batch = -1
        self.max_spec_tree_token_num = self.max_sequence_length = -1
        self._widths, self._num_ssms, self._eos = [], 0, []

    # ---- configuration (request_manager.cc:117-171, 181-263) ----
    def set_max_requests_per_batch(self, n):
        self.max_requests_per_batch = n

    def set_max_tokens_per_batch(self, n):
        self.max_tokens_per_batch = n

    def set_max_spec_tree_token_num(self, n):
        self.max_spec_tree_token_num = n

    def set_max_sequence_length(self, n):
        self.max_sequence_length = n

    def get_max_verify_tokens_per_batch(self):
        return self.max_tokens_per_batch + self.max_spec_tree_token_num * self.max_requests_per_batch

    def push_spec_infer_tree_width(self, tree_width):
        assert self._h is None, "configure before the first request"
        self._widths.append(tree_width)

    def register_ssm_model(self, model=None):
        assert self._h is None, "configure before the first request"
        self._num_ssms += 1
        return self._num_ssms - 1

    def register_tokenizer(self, eos_token_ids=()):
        self._eos = list(eos_token_ids)

    def _handle(self):
        if self._h is None:
            d = _lib.RMDesc(self.max_requests_per_batch, self.max_tokens_per_batch, max(self.max_spec_tree_token_num, 0),
                            self.max_sequence_length, self._num_ssms, len(self._eos))
            for i, e in enumerate(self._eos):
                d.eos_token_ids[i] = e
            h = C.c_void_p()
            _lib.check(_lib.load().ffb_rm_create(C.byref(d), C.byref(h)), "ffb_rm_create")
            for w in self._widths:
                _lib.check(_lib.load().ffb_rm_push_spec_infer_tree_width(h, w), "ffb_rm_push_spec_infer_tree_width")
            self._h = h
        return self._h

    def close(self):
        if self._h is not None:
            _lib.load().ffb_rm_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # noqa: BLE001
            pass

    # ---- requests ----
    def register_new_request(self, tokens, max_length=-1, max_new_tokens=-1):
        arr = (C.c_int * len(tokens))(*[int(t) for t in tokens])
        guid = C.c_size_t(0)
        _lib.check(_lib.load().ffb_rm_register_request(self._handle(), arr, len(tokens), max_length, max_new_tokens,
                                                       C.byref(guid)), "ffb_rm_register_request")
        return guid.value

    def request_status(self, guid):
        s = C.c_int(0)
        _lib.check(_lib.load().ffb_rm_request_status(self._handle(), guid, C.byref(s)), "ffb_rm_request_status")
        return s.value

    def request_tokens(self, guid):
        n = C.c_int(0)
        _lib.check(_lib.load().ffb_rm_request_tokens(self._handle(), guid, None, 0, C.byref(n)), "ffb_rm_request_tokens")
        arr = (C.c_int * max(1, n.value))()
        _lib.check(_lib.load().ffb_rm_request_tokens(self._handle(), guid, arr, n.value, C.byref(n)), "ffb_rm_request_tokens")
        return list(arr[:n.value])

    def request_steps(self, guid):
        a, b = C.c_int(0), C.c_int(0)
        _lib.check(_lib.load().ffb_rm_request_steps(self._handle(), guid, C.byref(a), C.byref(b)), "ffb_rm_request_steps")
        return a.value, b.value

    def num_pending(self):
        return _lib.load().ffb_rm_num_pending(self._handle())

    def get_num_processed_requests(self):
        return _lib.load().ffb_rm_num_processed_requests(self._handle())

    # ---- the batch builders ----
    def prepare_next_batch(self, old_bc, result):
        new_bc = B.BatchConfig()
        _lib.check(_lib.load().ffb_rm_prepare_next_batch(self._handle(), C.addressof(old_bc), C.addressof(result),
                                                         C.addressof(new_bc)), "ffb_rm_prepare_next_batch")
        return new_bc

    def prepare_next_batch_init(self, old_bc, result, model_id=0):
        new_bc = B.BeamSearchBatchConfig()
        _lib.check(_lib.load().ffb_rm_prepare_next_batch_init(self._handle(), C.addressof(old_bc), C.addressof(result),
                                                              model_id, C.addressof(new_bc)),
                   "ffb_rm_prepare_next_batch_init")
        return new_bc

    def prepare_next_batch_beam(self, old_bc, result):
        new_bc = B.BeamSearchBatchConfig()
        _lib.check(_lib.load().ffb_rm_prepare_next_batch_beam(self._handle(), C.addressof(old_bc), C.addressof(result),
                                                              C.addressof(new_bc)), "ffb_rm_prepare_next_batch_beam")
        return new_bc

    def prepare_next_batch_verify(self, old_batches):
        new_bc = B.TreeVerifyBatchConfig()
        ptrs = (C.c_void_p * len(old_batches))(*[C.addressof(b) for b in old_batches])
        _lib.check(_lib.load().ffb_rm_prepare_next_batch_verify(self._handle(), ptrs, len(old_batches),
                                                                C.addressof(new_bc)), "ffb_rm_prepare_next_batch_verify")
        return new_bc

    # ---- the serving loops, synchronous (request_manager.cc:3012-3081, 3083-3173 without the Legion future pipeline) ----
    def serve_incr_decoding(self, llm_step, on_batch=None, max_steps=1 << 30):
        """llm_step(BatchConfig) -> InferenceResult.  Runs until no request is active or queued."""
        bc, ir = B.BatchConfig().reset(), B.InferenceResult()
        for _ in range(max_steps):
            bc = self.prepare_next_batch(bc, ir)
            if on_batch:
                on_batch("incr", bc)
            if bc.num_active_requests() == 0 and self.num_pending() == 0:
                break
            ir = llm_step(bc)
        return bc

    def serve_spec_infer(self, llm_step, ssm_step, on_batch=None, max_rounds=1 << 30):
        """llm_step(TreeVerifyBatchConfig) -> InferenceResult; ssm_step(BeamSearchBatchConfig) -> BeamInferenceResult.
        One round = init, MAX_BEAM_DEPTH x (SSM step, beam), verify, LLM step."""
        tree_bc, tree_ir = B.TreeVerifyBatchConfig().reset(), B.InferenceResult()
        for _ in range(max_rounds):
            beam_bc = self.prepare_next_batch_init(tree_bc, tree_ir, 0)
            if on_batch:
                on_batch("init", beam_bc)
            if beam_bc.num_active_requests() == 0 and self.num_pending() == 0:
                break
            for _depth in range(B.MAX_BEAM_DEPTH):
                beam_ir = ssm_step(beam_bc)
                beam_bc = self.prepare_next_batch_beam(beam_bc, beam_ir)
                if on_batch:
                    on_batch("beam", beam_bc)
            tree_bc = self.prepare_next_batch_verify([beam_bc])
            if on_batch:
                on_batch("verify", tree_bc)
            tree_ir = llm_step(tree_bc)
        return tree_bc
