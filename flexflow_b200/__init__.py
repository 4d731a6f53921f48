"""flexflow_b200 — B200-native (sm_100a) implementation of FlexFlow Serve's incremental / speculative /
tree-verify attention hot path behind FlexFlow's own operator interface.  See DESIGN.md."""
from . import batch_config  # noqa: F401
from .batch_config import (  # noqa: F401
    BatchConfig, BeamSearchBatchConfig, TreeVerifyBatchConfig,
    DT_BF16, DT_FLOAT, DT_HALF, DT_INT4, DT_INT8, DT_NONE,
    INC_DECODING_MODE, BEAM_SEARCH_MODE, TREE_VERIFY_MODE,
)

__all__ = ["batch_config", "lib", "ops"]
