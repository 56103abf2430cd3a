"""Public op entry points with per-call dispatch between the sm_100a kernels and the oracles.

Counterparts of ``cs336_basics/nn_utils.py:4-30`` and ``model.py:400-432,526-527``.
"""

from __future__ import annotations

import torch
from torch import Tensor

from .. import _native
from . import functional as F_ref

def softmax(x: Tensor, dim: int = -1) -> Tensor:
    """Max-subtracted softmax (``nn_utils.py:4-7``).  CUDA fp32 / bf16 tensors reduced over their last dimension run
    the row kernel of ``csrc/softmax.cu`` (one pass over the row, registers in between); everything else the oracle."""
    if (dim == -1 or dim == x.dim() - 1) and x.dim() >= 1 and x.dtype in (torch.float32, torch.bfloat16) and \
            _native.use_native(x):
        from . import native

        return native.softmax_lastdim(x)
    return F_ref.softmax(x, dim)


log_softmax = F_ref.log_softmax
silu = F_ref.silu
scaled_dot_product_attention = F_ref.sdpa_ref


def cross_entropy(inputs: Tensor, targets: Tensor) -> Tensor:
    """Mean token cross-entropy.  CUDA: one fused online-logsumexp kernel (+ fused backward)."""
    if _native.use_native(inputs, targets) and inputs.dtype in (torch.bfloat16, torch.float32):
        from . import native

        return native.cross_entropy(inputs, targets)
    return F_ref.cross_entropy_ref(inputs, targets)


def clip_gradient(parameters, max_norm: float, eps: float = 1e-6) -> Tensor:
    """Global-L2 gradient clip.  CUDA fp32 grads: multi-tensor norm + scale kernels (2 launches)."""
    params = [p for p in parameters if p.grad is not None]
    if params and all(p.grad.is_cuda and p.grad.dtype == torch.float32 and p.grad.is_contiguous() for p in params) \
            and not _native.force_torch():  # a scaled .contiguous() copy would leave the real gradient unclipped
        from . import native

        return native.clip_grad_norm_([p.grad for p in params], max_norm, eps)
    return F_ref.clip_gradient_ref(params, max_norm, eps)
