"""Transformer building blocks, state-dict compatible with the course library the reference
drives (``cs336-basics/cs336_basics/model.py:22-150,330-527``).

Every module has two execution paths chosen per call by :func:`native_active`:

* **oracle path** (CPU tensors, fp32 CUDA without autocast, or ``B200SYS_FORCE_TORCH=1``):
  plain PyTorch ops with the reference's exact arithmetic;
* **native path** (CUDA tensors under ``torch.autocast('cuda', bfloat16)``): the hand-written
  sm_100a kernels — tcgen05 GEMMs on cached bf16 weight shadows (q/k/v and w1/w3 share one
  concatenated shadow so each pair is ONE GEMM), fused RMSNorm(+residual add), in-place
  RoPE + tcgen05 FlashAttention-2, SwiGLU gate.

Parameter names, shapes, init and ``parameters()`` order are unchanged, so checkpoints
(``model_config.json`` + ``model.pt``) load either way.
"""

from __future__ import annotations

import os

import math

import torch
import torch.nn as nn
from torch import Tensor

from .. import _native
from ..ops import functional as F_ref


def native_active(x: Tensor) -> bool:
    """Native kernels run for CUDA tensors under bf16 autocast (or bf16 activations)."""
    if not x.is_cuda or _native.force_torch():
        return False
    if torch.is_autocast_enabled("cuda") and torch.get_autocast_dtype("cuda") == torch.bfloat16:
        _native.lib()
        return True
    if x.dtype == torch.bfloat16:
        _native.lib()
        return True
    return False


class Linear(nn.Module):
    """Bias-free linear, ``weight (d_out, d_in)``, trunc-normal σ=√(2/(d_in+d_out)) clipped ±3σ
    (model.py:22-44)."""

    def __init__(self, d_in: int, d_out: int, device=None, dtype=None):
        super().__init__()
        std = math.sqrt(2.0 / (d_in + d_out))
        w = torch.empty(d_out, d_in, device=device, dtype=dtype)
        nn.init.trunc_normal_(w, std=std, a=-3 * std, b=3 * std)
        self.weight = nn.Parameter(w)

    def forward(self, x: Tensor) -> Tensor:
        if native_active(x):
            from ..ops import native

            return native.linear(x, (self.weight,))
        return torch.nn.functional.linear(x, self.weight)

    def extra_repr(self):
        return f"d_out={self.weight.shape[0]}, d_in={self.weight.shape[1]}"


class Embedding(nn.Module):
    """``weight (vocab, d_model)`` trunc-normal σ=1 ±3; forward is a row gather (model.py:47-60)."""

    def __init__(self, vocab_size: int, d_model: int, device=None, dtype=None):
        super().__init__()
        w = torch.empty(vocab_size, d_model, device=device, dtype=dtype)
        nn.init.trunc_normal_(w, std=1.0, a=-3.0, b=3.0)
        self.weight = nn.Parameter(w)

    def forward(self, token_ids: Tensor) -> Tensor:
        if token_ids.is_cuda and not _native.force_torch() and self.weight.dtype == torch.float32:
            from ..ops import native

            return native.embedding(token_ids, self.weight)
        return self.weight[token_ids]

    def extra_repr(self):
        return f"vocab_size={self.weight.shape[0]}, d={self.weight.shape[1]}"


class RMSNorm(nn.Module):
    """fp32-internal RMSNorm, result in the input dtype (model.py:63-110)."""

    def __init__(self, hidden_size: int, eps: float = 1e-5, device=None):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size, device=device))
        self.eps = eps

    def forward(self, x: Tensor, out_dtype: torch.dtype | None = None) -> Tensor:
        if x.is_cuda and not _native.force_torch() and x.dtype in (torch.float32, torch.bfloat16):
            from ..ops import native

            return native.rmsnorm(x, self.weight, self.eps, out_dtype or x.dtype)
        return F_ref.rmsnorm_ref(x, self.weight, self.eps, out_dtype)

    def extra_repr(self):
        return f"hidden_size={self.weight.shape[0]}, eps={self.eps}"


class RotaryEmbedding(nn.Module):
    """RoPE on interleaved pairs with a non-persistent ``(2, ctx, dim/2)`` cos/sin cache
    (model.py:113-150).  Not part of the state dict."""

    def __init__(self, context_length: int, dim: int, theta: float = 10000.0):
        super().__init__()
        self.register_buffer("_freq_cis_cache", F_ref.rope_cache(context_length, dim, theta), persistent=False)

    def forward(self, x: Tensor, pos_ids: Tensor | None = None) -> Tensor:
        return F_ref.rope_ref(x, self._freq_cis_cache, pos_ids)

    def extra_repr(self):
        return f"context_length={self._freq_cis_cache.shape[1]}, dim/2={self._freq_cis_cache.shape[2]}"


class SwiGLU(nn.Module):
    """``w2(silu(w1 x) * w3 x)`` (model.py:389-397)."""

    def __init__(self, d_model: int, d_ff: int):
        super().__init__()
        self.w1 = Linear(d_model, d_ff)
        self.w2 = Linear(d_ff, d_model)
        self.w3 = Linear(d_model, d_ff)

    def forward(self, x: Tensor) -> Tensor:
        if native_active(x):
            from ..ops import native

            if not torch.is_grad_enabled():  # serving / eval: the gate runs in the GEMM epilogue
                h = native.swiglu_linear_nograd(x, self.w1.weight, self.w3.weight)
                return native.linear(h, (self.w2.weight,))
            mode = os.environ.get("B200SYS_FUSE_SWIGLU", "1")  # 0: separate gate kernels, 1: forward epilogue, 2: both
            if mode != "0":
                native.SwiGLUFFNFn.FUSE_BACKWARD = mode == "2"
                y = native.swiglu_ffn(x, self.w1.weight, self.w3.weight, self.w2.weight)  # gate in the GEMM epilogue(s)
                if y is not None:
                    return y
            ab = native.linear(x, (self.w1.weight, self.w3.weight))  # one GEMM → [.., 2*d_ff]
            return native.linear(native.swiglu(ab), (self.w2.weight,))
        return self.w2(F_ref.silu(self.w1(x)) * self.w3(x))


class CausalMultiHeadSelfAttention(nn.Module):
    """Causal MHA with separate q/k/v/output projections and RoPE on Q, K (model.py:435-524)."""

    def __init__(self, d_model: int, num_heads: int, positional_encoder: RotaryEmbedding):
        super().__init__()
        assert d_model % num_heads == 0
        self.d_model = d_model
        self.num_heads = num_heads
        self.d_k = d_model // num_heads
        self.d_v = self.d_k
        self.q_proj = Linear(d_model, num_heads * self.d_k)
        self.k_proj = Linear(d_model, num_heads * self.d_k)
        self.v_proj = Linear(d_model, num_heads * self.d_v)
        self.output_proj = Linear(num_heads * self.d_v, d_model)
        self.positional_encoder = positional_encoder

    def forward(self, x: Tensor, token_positions: Tensor | None = None) -> Tensor:
        *b, S, d_model = x.shape
        assert d_model == self.d_model
        H, dh = self.num_heads, self.d_k
        if native_active(x) and token_positions is None:
            from ..ops import native

            cache = self.positional_encoder._freq_cis_cache
            # one GEMM for q,k,v; its epilogue applies RoPE to the q and k columns
            qkv = native.linear(x, (self.q_proj.weight, self.k_proj.weight, self.v_proj.weight),
                                rope=(cache, S, dh, 2 * H * dh))
            o = native.attention_core(qkv.view(-1, S, 3, H, dh), cache, rope_done=True)
            return native.linear(o.view(*b, S, d_model), (self.output_proj.weight,))
        q, k, v = self.q_proj(x), self.k_proj(x), self.v_proj(x)
        q, k, v = (t.view(*b, S, H, dh).transpose(-2, -3) for t in (q, k, v))  # (..., H, S, dh)
        if token_positions is not None:
            token_positions = token_positions.unsqueeze(-2)  # broadcast over heads
        q = self.positional_encoder(q, token_positions)
        k = self.positional_encoder(k, token_positions)
        seq = torch.arange(S, device=x.device)
        causal = seq[:, None] >= seq[None, :]
        o = F_ref.sdpa_ref(q, k, v, causal)
        o = o.transpose(-2, -3).reshape(*b, S, d_model)
        return self.output_proj(o)


class TransformerBlock(nn.Module):
    """Pre-norm block; submodules registered as ``attn, ffn, ln1, ln2`` (model.py:330-386) so
    ``parameters()`` order matches the reference."""

    def __init__(self, d_model: int, num_heads: int, d_ff: int, positional_encoder: RotaryEmbedding):
        super().__init__()
        self.attn = CausalMultiHeadSelfAttention(d_model, num_heads, positional_encoder)
        self.ffn = SwiGLU(d_model, d_ff)
        self.ln1 = RMSNorm(d_model)
        self.ln2 = RMSNorm(d_model)

    def forward(self, x: Tensor) -> Tensor:
        if native_active(x) and x.dtype == torch.float32:
            x, y = self.forward_native(x, None)
            from ..ops import native

            return native.residual_add(x, y)
        x = x + self.attn(self.ln1(x))
        return x + self.ffn(self.ln2(x))

    def forward_native(self, x: Tensor, y_pending: Tensor | None):
        """Native path with the residual add deferred into the next RMSNorm.

        Takes the fp32 residual stream ``x`` and an optional not-yet-added bf16 branch output
        ``y_pending``; returns ``(x', y')`` with the same meaning (``x' + y'`` is the block output).
        """
        from ..ops import native

        cd = torch.bfloat16
        if y_pending is None:
            h = native.rmsnorm(x, self.ln1.weight, self.ln1.eps, cd)
        else:
            x, h = native.add_rmsnorm(x, y_pending, self.ln1.weight, self.ln1.eps)
        y1 = self.attn(h)
        x, h = native.add_rmsnorm(x, y1, self.ln2.weight, self.ln2.eps)
        y2 = self.ffn(h)
        return x, y2
