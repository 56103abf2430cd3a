"""NVTX-annotated model for timeline profiling.

The reference keeps a hand-annotated copy of its Transformer (``cs336_systems/transformer_annotated.py:1-185``,
which no longer imports) and monkey-patches it in for ``nsys`` runs.  Here any model is annotated in place
with forward pre/post hooks: one NVTX range per interesting submodule (embedding, each block, attention,
feed-forward, norms, projections, lm_head) and a matching range around its backward, so ``nsys``/``ncu``
timelines are readable without a second copy of the model.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..utils import nvtx as _nvtx  # noqa: F401  (keeps the no-CUDA fallback importable)
from .layers import (CausalMultiHeadSelfAttention, Embedding, Linear, RMSNorm, SwiGLU, TransformerBlock)

_KINDS = (TransformerBlock, CausalMultiHeadSelfAttention, SwiGLU, RMSNorm, Embedding, Linear)


def annotate(model: nn.Module, kinds=_KINDS) -> list:
    """Attach NVTX push/pop hooks to every submodule of the given kinds; returns the hook handles."""
    handles = []
    if not torch.cuda.is_available():
        return handles
    for name, mod in model.named_modules():
        if not name or not isinstance(mod, kinds):
            continue
        label = f"{type(mod).__name__}:{name}"

        def pre(_m, _inp, label=label):
            torch.cuda.nvtx.range_push(label)

        def post(_m, _inp, _out):
            torch.cuda.nvtx.range_pop()

        def bpre(_m, _g, label=label):
            torch.cuda.nvtx.range_push(label + ":bwd")

        def bpost(_m, _gi, _go):
            torch.cuda.nvtx.range_pop()

        handles += [mod.register_forward_pre_hook(pre), mod.register_forward_hook(post),
                    mod.register_full_backward_pre_hook(bpre), mod.register_full_backward_hook(bpost)]
    return handles


def remove(handles) -> None:
    for h in handles:
        h.remove()
