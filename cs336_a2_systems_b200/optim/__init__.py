"""Optimizers: AdamW with course semantics (+ fused sm_100a multi-tensor kernel), cosine LR schedule."""
from .adamw import AdamW, get_cosine_lr

__all__ = ["AdamW", "get_cosine_lr"]
