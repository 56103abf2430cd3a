"""Pure-PyTorch definitions of every op in the library.

These are (a) the CPU execution path, (b) the fp32 oracles the GPU numerics tests compare
the sm_100a kernels against.  Semantics follow the course library the reference drives
(``cs336-basics/cs336_basics/nn_utils.py:4-30``, ``model.py:101-107,135-147,397,400-432``),
restated without ``einx``.
"""

from __future__ import annotations

import math

import torch
from torch import Tensor


def softmax(x: Tensor, dim: int = -1) -> Tensor:
    """Max-subtracted softmax (nn_utils.py:4-7)."""
    z = x - x.amax(dim=dim, keepdim=True)
    e = torch.exp(z)
    return e / e.sum(dim=dim, keepdim=True)


def log_softmax(x: Tensor, dim: int = -1) -> Tensor:
    z = x - x.amax(dim=dim, keepdim=True)
    return z - torch.log(torch.exp(z).sum(dim=dim, keepdim=True))


def cross_entropy_ref(logits: Tensor, targets: Tensor) -> Tensor:
    """Mean over all tokens of the negative gathered log-softmax (nn_utils.py:15-17)."""
    nls = -log_softmax(logits.float() if logits.dtype != torch.float64 else logits)
    return torch.gather(nls, -1, targets.unsqueeze(-1)).mean()


def silu(x: Tensor) -> Tensor:
    return x * torch.sigmoid(x)


def swiglu_ref(a: Tensor, b: Tensor) -> Tensor:
    """``silu(a) * b`` — the SwiGLU gate (model.py:397)."""
    return silu(a) * b


def rmsnorm_ref(x: Tensor, weight: Tensor, eps: float = 1e-5, out_dtype: torch.dtype | None = None) -> Tensor:
    """fp32-internal RMSNorm (model.py:101-107); result cast to ``out_dtype`` (default: x.dtype)."""
    out_dtype = out_dtype or x.dtype
    xf = x.to(torch.float32)
    rstd = torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (weight.to(torch.float32) * (xf * rstd)).to(out_dtype)


def rope_cache(context_length: int, dim: int, theta: float = 10000.0, device=None) -> Tensor:
    """``(2, ctx, dim/2)`` cos/sin table of ``t * theta^(-2i/dim)`` (model.py:121-132)."""
    assert dim % 2 == 0
    d = torch.arange(0, dim, 2, device=device, dtype=torch.float32) / dim
    freqs = theta ** -d
    t = torch.arange(context_length, device=device, dtype=torch.float32)
    ang = t[:, None] * freqs[None, :]
    return torch.stack((torch.cos(ang), torch.sin(ang)))


def rope_ref(x: Tensor, cache: Tensor, pos: Tensor | None = None, inverse: bool = False) -> Tensor:
    """Rotate interleaved pairs ``(x[2i], x[2i+1])`` of the last dim (model.py:135-147).

    ``x``: ``(..., seq, d)``; ``pos``: broadcastable to ``(..., seq)`` or None for arange.
    ``inverse`` applies the transposed rotation (the backward of the op).
    """
    seq = x.shape[-2]
    if pos is None:
        cos, sin = cache[0, :seq], cache[1, :seq]
    else:
        cos, sin = cache[0][pos], cache[1][pos]
    if inverse:
        sin = -sin
    cos, sin = cos.to(x.dtype), sin.to(x.dtype)
    x1, x2 = x[..., 0::2], x[..., 1::2]
    r1 = cos * x1 - sin * x2
    r2 = sin * x1 + cos * x2
    return torch.stack((r1, r2), dim=-1).flatten(-2)


def sdpa_ref(q: Tensor, k: Tensor, v: Tensor, mask: Tensor | None = None) -> Tensor:
    """Materialised scaled-dot-product attention (model.py:400-432)."""
    s = (q @ k.transpose(-1, -2)) / math.sqrt(k.shape[-1])
    if mask is not None:
        s = torch.where(mask, s, float("-inf"))
    return softmax(s, dim=-1) @ v


def attention_and_lse_ref(q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False):
    """Oracle used by the attention tests: output and per-row logsumexp of the scaled scores."""
    d = q.shape[-1]
    s = torch.einsum("...qd,...kd->...qk", q.float(), k.float()) / math.sqrt(d)
    if is_causal:
        nq, nk = s.shape[-2:]
        m = torch.arange(nq, device=s.device)[:, None] >= torch.arange(nk, device=s.device)[None, :]
        s = torch.where(m, s, torch.full_like(s, -1e6))
    p = torch.softmax(s, dim=-1)
    return p @ v.float(), torch.logsumexp(s, dim=-1)


def clip_gradient_ref(parameters, max_norm: float, eps: float = 1e-6) -> Tensor:
    """Global-L2 clip with in-place scaling (nn_utils.py:20-30). Returns the pre-clip norm."""
    grads = [p.grad for p in parameters if p.grad is not None]
    if not grads:
        return torch.tensor(0.0)
    total = torch.zeros((), dtype=torch.float32, device=grads[0].device)
    for g in grads:
        total = total + g.detach().float().pow(2).sum()
    norm = total.sqrt()
    coef = torch.clamp(max_norm / (norm + eps), max=1.0)
    for g in grads:
        g.mul_(coef.to(g.dtype))
    return norm
