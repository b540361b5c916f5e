"""Custom ops.  CUDA tensors -> hand-written sm_100a kernels (``csrc/``) through the C ABI;
CPU tensors -> the fp32 PyTorch restatement in :mod:`.reference` (tests / plumbing only)."""
from . import reference  # noqa: F401
from .flat import (masked_sgd_step, weighted_aggregate, saliency_accumulate, global_topk_mask,  # noqa: F401
                   count_nonzero, hamming_distance, squared_distance, prox_pull, grad_norm, prunable_flags, gather_rows_u8, mask_select_k)
