"""sparkinfer_b200 -- B200-native (sm_100a) kernels for SparkInfer's activation-sparse FFN hot path.

The product is the C-ABI shared library (include/spif_b200.h, sparkinfer_b200/libspif_b200.so, sources in
sparkinfer_b200/csrc/).  This package is the thin host-side mirror of the reference's ggml operator surface used by
the tests and bench.py: torch supplies device memory, streams and torch.distributed; every op calls the CUDA
library and raises if it is missing -- there is no CPU or eager fallback.
"""
from . import _lib  # noqa: F401
from .ops import (  # noqa: F401
    ARCH_PROSPARSE, ARCH_RELU_GATED, ARCH_RELU_UP, ARCH_SILU_GATED, ActiveList, FLAG_CPU_ROUNDING,
    FLAG_MASK_BY_ROW, FLAG_MASK_STABLE, FLAG_PER_TOKEN, GGML_TYPE_BF16, GGML_TYPE_F16, GGML_TYPE_Q4_0, GGML_TYPE_Q8_0, Workspace, axpy_sparse,
    ShardComm, dfr_update, initial_placement, rebalance_plan, fatrelu, mask_compact, mul_mat_sparse, predictor, row_size, shard_allreduce, shifted_step, sparse_ffn,
    topk_mask, mask_to_list, sparse_ffn_lookahead)

from .balancer import ShardBalancer  # noqa: F401,E402

__all__ = [n for n in dir() if not n.startswith("_")]
