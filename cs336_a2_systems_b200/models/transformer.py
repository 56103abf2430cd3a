"""``TransformerLM`` — drop-in for ``BasicsTransformerLM`` (``cs336_basics/model.py:153-327``).

Same constructor, ``config`` dict, member names (``token_embeddings``, ``positional_encoder``,
``layers``, ``ln_final``, ``lm_head``), state-dict keys, ``get_num_params``, ``generate`` and
``from_pretrained`` (``model_config.json`` + ``model.pt``, ``_orig_mod.`` prefix tolerated).
The forward runs the native sm_100a path under bf16 autocast on CUDA (see ``layers.py``).
"""

from __future__ import annotations

import json
import logging
import os

import torch
import torch.nn as nn
from torch import Tensor

from ..ops import functional as F_ref
from .layers import Embedding, Linear, RMSNorm, RotaryEmbedding, TransformerBlock, native_active

logger = logging.getLogger(__name__)

#: Model presets of the course handout (reference ``benchmark.py:247-259``): d_model, d_ff, layers, heads.
MODEL_CONFIGS: dict[str, dict[str, int]] = {
    "small": dict(d_model=768, d_ff=3072, num_layers=12, num_heads=12),
    "medium": dict(d_model=1024, d_ff=4096, num_layers=24, num_heads=16),
    "large": dict(d_model=1280, d_ff=5120, num_layers=36, num_heads=20),
    "xl": dict(d_model=1600, d_ff=6400, num_layers=48, num_heads=25),
    "2.7b": dict(d_model=2560, d_ff=10240, num_layers=32, num_heads=32),
}


def get_model_config(size: str) -> dict[str, int]:
    try:
        return dict(MODEL_CONFIGS[size.lower()])
    except KeyError as e:
        raise ValueError(f"unknown model size {size!r}; choose from {sorted(MODEL_CONFIGS)}") from e


def build_lm(size: str = "small", vocab_size: int = 10000, context_length: int = 512,
             rope_theta: float = 10000.0, device=None) -> "TransformerLM":
    cfg = get_model_config(size)
    if device is not None and torch.device(device).type in ("cuda", "meta"):
        # initialise on the GPU: trunc_normal_ of 2B parameters on the host is minutes ("meta": shapes only)
        with torch.device(device):
            return TransformerLM(vocab_size=vocab_size, context_length=context_length,
                                 rope_theta=rope_theta, **cfg)
    m = TransformerLM(vocab_size=vocab_size, context_length=context_length, rope_theta=rope_theta, **cfg)
    return m.to(device) if device is not None else m


class TransformerLM(nn.Module):
    def __init__(self, vocab_size: int, context_length: int, d_model: int, num_layers: int,
                 num_heads: int, d_ff: int, rope_theta: float = 10000.0):
        self.config = dict(vocab_size=vocab_size, context_length=context_length, d_model=d_model,
                           num_layers=num_layers, num_heads=num_heads, d_ff=d_ff, rope_theta=rope_theta)
        super().__init__()
        self.vocab_size = vocab_size
        self.context_length = context_length
        self.d_model = d_model
        self.token_embeddings = Embedding(vocab_size, d_model)
        self.positional_encoder = RotaryEmbedding(context_length, d_model // num_heads, rope_theta)
        self.layers = nn.ModuleList(
            [TransformerBlock(d_model, num_heads, d_ff, self.positional_encoder) for _ in range(num_layers)]
        )
        self.ln_final = RMSNorm(d_model)
        self.lm_head = Linear(d_model, vocab_size)
        logger.info("number of non-embedding parameters: %.2fM", self.get_num_params() / 1e6)

    def get_num_params(self, non_embedding: bool = True) -> int:
        n = sum(p.numel() for p in self.parameters())
        if non_embedding:
            n -= self.lm_head.weight.numel()
        return n

    def forward(self, x: Tensor) -> Tensor:
        """``x``: int64 ``(B, S)`` token ids → logits ``(B, S, vocab)``."""
        h = self.token_embeddings(x)
        if native_active(h) and h.dtype == torch.float32:
            from ..ops import native

            y = None
            for layer in self.layers:
                h, y = layer.forward_native(h, y)
            if y is None:
                hn = native.rmsnorm(h, self.ln_final.weight, self.ln_final.eps, torch.bfloat16)
            else:
                _, hn = native.add_rmsnorm(h, y, self.ln_final.weight, self.ln_final.eps)
            return self.lm_head(hn)
        for layer in self.layers:
            h = layer(h)
        return self.lm_head(self.ln_final(h))

    @torch.no_grad()
    def generate(self, x: Tensor, max_new_tokens: int, temperature: float = 1.0,
                 top_k: int | None = None, eos_token_id: int | None = None) -> Tensor:
        """Ancestral sampling with temperature / top-k (model.py:255-310; here top-k really masks —
        the reference's out-of-place ``masked_fill`` at ``:302`` makes it a no-op)."""
        if x.dim() == 1:
            x = x.unsqueeze(0)
        n0 = x.size(-1)
        for _ in range(max_new_tokens):
            ctx = x[:, -self.context_length:] if x.size(1) > self.context_length else x
            logits = self.forward(ctx)[:, -1].float() / temperature
            if top_k:
                kth = torch.topk(logits, min(top_k, logits.size(-1))).values[:, -1:]
                logits = logits.masked_fill(logits < kth, float("-inf"))
            nxt = torch.multinomial(F_ref.softmax(logits, dim=-1), 1)
            if eos_token_id is not None and nxt.numel() == 1 and nxt.item() == eos_token_id:
                break
            x = torch.cat((x, nxt), dim=-1)
        return x[:, n0:]

    # -- checkpoint format (model.py:312-327) ------------------------------------------------
    def save_pretrained(self, path: str) -> None:
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, "model_config.json"), "w") as f:
            json.dump(self.config, f)
        torch.save({k: v.detach().cpu() for k, v in self.state_dict().items()}, os.path.join(path, "model.pt"))

    @classmethod
    def from_pretrained(cls, pretrained_model_path: str, map_location="cpu") -> "TransformerLM":
        with open(os.path.join(pretrained_model_path, "model_config.json")) as f:
            config = json.load(f)
        model = cls(**config)
        sd = torch.load(os.path.join(pretrained_model_path, "model.pt"), map_location=map_location)
        prefix = "_orig_mod."
        sd = {(k[len(prefix):] if k.startswith(prefix) else k): v for k, v in sd.items()}
        model.load_state_dict(sd)
        return model


#: alias so code written against the course library keeps working
BasicsTransformerLM = TransformerLM
