"""Op library: PyTorch oracles (``functional``), FlashAttention-2 Functions (``attention``),
ctypes bindings of the sm_100a kernels (``kernels``) and their autograd wrappers (``native``)."""
from . import functional
from .attention import FlashAttentionSM100, FlashAttentionTorch, flash_attention
from .api import clip_gradient, cross_entropy, log_softmax, scaled_dot_product_attention, silu, softmax

__all__ = ["functional", "FlashAttentionTorch", "FlashAttentionSM100", "flash_attention", "softmax",
           "log_softmax", "cross_entropy", "clip_gradient", "silu", "scaled_dot_product_attention"]
