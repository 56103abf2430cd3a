"""The reference's public names, for code written against ``kkaitlyn111/cs336-a2-systems``.

A user of the reference imports ``FlashAttentionTriton`` / ``FlashAttentionTorch``
(``cs336_systems/flash_attention.py:8,85``), ``DDP`` / ``DDP_Bucketed`` / ``ShardedStateOptimizer``
(``cs336_systems/ddp_bucketed_overlapped_sharded.py:217,251,322``), ``get_model_config`` /
``run_simple_benchmark`` / ``run_memory_foward_fullstep`` (``cs336_systems/benchmark.py:27,175,247``) and, from the
course library, ``BasicsTransformerLM``, ``AdamW``, ``get_cosine_lr``, ``cross_entropy``, ``clip_gradient``,
``get_batch`` (``cs336_basics/model.py:153``, ``optimizer.py:9,30``, ``nn_utils.py:9,20``, ``data.py:10``).
Switching is a one-line import change::

    from cs336_a2_systems_b200.compat import FlashAttentionTriton, DDP_Bucketed, ShardedStateOptimizer

These are the same objects as the native names (no wrappers, no second code path).
"""
from .bench.lm import run_simple_benchmark
from .bench.memory import run_memory_forward_fullstep
from .models import MODEL_CONFIGS, TransformerLM, get_model_config
from .models.layers import (CausalMultiHeadSelfAttention, Embedding, Linear, RMSNorm, RotaryEmbedding, SwiGLU,
                            TransformerBlock)
from .ops import (FlashAttentionSM100, FlashAttentionTorch, clip_gradient, cross_entropy, log_softmax,
                  scaled_dot_product_attention, silu, softmax)
from .optim import AdamW, get_cosine_lr
from .parallel import DDPBucketed, DDPIndividualParameters, ShardedOptimizer
from .train.data import get_batch

BasicsTransformerLM = TransformerLM
FlashAttentionTriton = FlashAttentionSM100  # the hand-written GPU kernel behind the same autograd.Function contract
DDP = DDPIndividualParameters
DDP_Bucketed = DDPBucketed
ShardedStateOptimizer = ShardedOptimizer
run_memory_foward_fullstep = run_memory_forward_fullstep  # the reference's spelling

__all__ = ["BasicsTransformerLM", "TransformerLM", "TransformerBlock", "CausalMultiHeadSelfAttention", "SwiGLU",
           "RMSNorm", "RotaryEmbedding", "Linear", "Embedding", "FlashAttentionTorch", "FlashAttentionTriton",
           "FlashAttentionSM100", "DDP", "DDP_Bucketed", "ShardedStateOptimizer", "AdamW", "get_cosine_lr",
           "cross_entropy", "clip_gradient", "softmax", "log_softmax", "silu", "scaled_dot_product_attention",
           "get_batch", "get_model_config", "MODEL_CONFIGS", "run_simple_benchmark", "run_memory_forward_fullstep",
           "run_memory_foward_fullstep"]
