"""Model family: the course Transformer LM and its building blocks."""
from .layers import (CausalMultiHeadSelfAttention, Embedding, Linear, RMSNorm, RotaryEmbedding, SwiGLU,
                     TransformerBlock)
from .transformer import (MODEL_CONFIGS, BasicsTransformerLM, TransformerLM, build_lm, get_model_config)

__all__ = ["Linear", "Embedding", "RMSNorm", "RotaryEmbedding", "SwiGLU", "CausalMultiHeadSelfAttention",
           "TransformerBlock", "TransformerLM", "BasicsTransformerLM", "MODEL_CONFIGS", "get_model_config",
           "build_lm"]
