"""FlashAttention-2 autograd Functions.

``FlashAttentionTorch``  — tiled online-softmax forward + tiled recompute backward in plain
    PyTorch (CPU-capable; no ``torch.compile``).  Counterpart of the reference's
    ``FlashAttentionTorch`` (``cs336_systems/flash_attention.py:8-83``) but, unlike it, the
    forward honours ``is_causal``.
``FlashAttentionSM100`` — forward and backward on hand-written tcgen05/TMEM/TMA kernels
    (``csrc/fa2_fwd_sm100.cu``, ``csrc/fa2_bwd_sm100.cu``).  Counterpart of
    ``FlashAttentionTriton`` + ``compiled_backward`` (``flash_attention.py:85-289``); the
    backward is tiled (O(N·d) memory) instead of materialising N×N matrices.

Contract shared by both (``tests/test_attention.py`` in the reference): call as
``apply(q, k, v, is_causal=False)``, inputs ``(B, N, D)`` (or 2-D), saved tensors are
``Q, K, V, O, L`` with exactly one tensor of shape ``(B, N_q)`` — the logsumexp ``L``.
"""

from __future__ import annotations

import math

import torch
from torch import Tensor

from .. import _native


def _as3d(t: Tensor):
    if t.dim() == 2:
        return t.unsqueeze(0), True
    if t.dim() > 3:
        return t.reshape(-1, t.shape[-2], t.shape[-1]), False
    return t, False


def _tiled_forward(q: Tensor, k: Tensor, v: Tensor, is_causal: bool, bq: int, bk: int):
    B, Nq, D = q.shape
    Nk = k.shape[1]
    scale = 1.0 / math.sqrt(D)
    qf, kf, vf = q.float(), k.float(), v.float()
    O = torch.empty(B, Nq, D, dtype=torch.float32, device=q.device)
    L = torch.empty(B, Nq, dtype=torch.float32, device=q.device)
    for i0 in range(0, Nq, bq):
        i1 = min(i0 + bq, Nq)
        qi = qf[:, i0:i1]
        m = torch.full((B, i1 - i0), float("-inf"), device=q.device)
        l = torch.zeros(B, i1 - i0, device=q.device)
        acc = torch.zeros(B, i1 - i0, D, device=q.device)
        k_end = min(Nk, i1) if is_causal else Nk  # causal: key tiles past the diagonal are skipped
        for j0 in range(0, k_end, bk):
            j1 = min(j0 + bk, Nk)
            s = torch.bmm(qi, kf[:, j0:j1].transpose(1, 2)) * scale
            if is_causal and j1 > i0 + 1:
                qi_idx = torch.arange(i0, i1, device=q.device)[:, None]
                kj_idx = torch.arange(j0, j1, device=q.device)[None, :]
                s = s.masked_fill(qi_idx < kj_idx, float("-inf"))
            m_new = torch.maximum(m, s.amax(dim=-1))
            p = torch.exp(s - m_new[..., None])
            alpha = torch.exp(m - m_new)
            l = alpha * l + p.sum(dim=-1)
            acc = alpha[..., None] * acc + torch.bmm(p, vf[:, j0:j1])
            m = m_new
        O[:, i0:i1] = acc / l[..., None]
        L[:, i0:i1] = m + torch.log(l)
    return O, L


def _tiled_backward(q, k, v, o, L, do, is_causal: bool, bq: int, bk: int):
    """FlashAttention-2 Algorithm 2: recompute P from L tile by tile; never builds N×N."""
    B, Nq, D = q.shape
    Nk = k.shape[1]
    scale = 1.0 / math.sqrt(D)
    qf, kf, vf, of, dof = q.float(), k.float(), v.float(), o.float(), do.float()
    delta = (of * dof).sum(dim=-1)  # (B, Nq)
    dq = torch.zeros_like(qf)
    dk = torch.zeros_like(kf)
    dv = torch.zeros_like(vf)
    for j0 in range(0, Nk, bk):
        j1 = min(j0 + bk, Nk)
        kj, vj = kf[:, j0:j1], vf[:, j0:j1]
        i_start = (j0 // bq) * bq if is_causal else 0
        for i0 in range(i_start, Nq, bq):
            i1 = min(i0 + bq, Nq)
            s = torch.bmm(qf[:, i0:i1], kj.transpose(1, 2)) * scale
            if is_causal and j1 > i0 + 1:
                qi_idx = torch.arange(i0, i1, device=q.device)[:, None]
                kj_idx = torch.arange(j0, j1, device=q.device)[None, :]
                s = s.masked_fill(qi_idx < kj_idx, float("-inf"))
            p = torch.exp(s - L[:, i0:i1, None])
            doi = dof[:, i0:i1]
            dv[:, j0:j1] += torch.bmm(p.transpose(1, 2), doi)
            dp = torch.bmm(doi, vj.transpose(1, 2))
            ds = p * (dp - delta[:, i0:i1, None]) * scale
            dq[:, i0:i1] += torch.bmm(ds, kj)
            dk[:, j0:j1] += torch.bmm(ds.transpose(1, 2), qf[:, i0:i1])
    return dq, dk, dv


class FlashAttentionTorch(torch.autograd.Function):
    """Tiled FA2 in PyTorch ops; tile sizes are class attributes (the reference fixes 16×16)."""

    BQ = 64
    BK = 64

    @staticmethod
    def forward(ctx, q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False):
        q3, squeeze = _as3d(q)
        k3, _ = _as3d(k)
        v3, _ = _as3d(v)
        O, L = _tiled_forward(q3, k3, v3, bool(is_causal), FlashAttentionTorch.BQ, FlashAttentionTorch.BK)
        O = O.to(q.dtype)
        ctx.save_for_backward(q3, k3, v3, O, L)
        ctx.is_causal = bool(is_causal)
        ctx.q_shape, ctx.k_shape, ctx.v_shape = q.shape, k.shape, v.shape
        return O.reshape(q.shape[:-1] + (v.shape[-1],)) if not squeeze else O[0]

    @staticmethod
    def backward(ctx, do: Tensor):
        q, k, v, o, L = ctx.saved_tensors
        do3, _ = _as3d(do)
        dq, dk, dv = _tiled_backward(q, k, v, o, L, do3, ctx.is_causal,
                                     FlashAttentionTorch.BQ, FlashAttentionTorch.BK)
        return (dq.to(q.dtype).reshape(ctx.q_shape), dk.to(k.dtype).reshape(ctx.k_shape),
                dv.to(v.dtype).reshape(ctx.v_shape), None)


class FlashAttentionSM100(torch.autograd.Function):
    """FA2 on the tcgen05 kernels.  ``(B, N, D)`` inputs, bf16 compute, fp32 softmax/accumulate.

    fp32 inputs are rounded to bf16 operands on the way in (the tensor cores take bf16) and
    the output is returned in the input dtype; ``L`` is always fp32.
    """

    @staticmethod
    def forward(ctx, q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False):
        from . import kernels

        if not q.is_cuda:
            raise RuntimeError("FlashAttentionSM100 needs CUDA tensors (use FlashAttentionTorch on CPU)")
        _native.lib()
        in_dtype = q.dtype
        q3, squeeze = _as3d(q)
        k3, _ = _as3d(k)
        v3, _ = _as3d(v)
        qb = q3.to(torch.bfloat16).contiguous()
        kb = k3.to(torch.bfloat16).contiguous()
        vb = v3.to(torch.bfloat16).contiguous()
        B, Nq, D = qb.shape
        # kernels take (B, H, N, D) with arbitrary batch/head/row strides; here H = 1
        o, L = kernels.fa2_fwd(qb.view(B, 1, Nq, D), kb.view(B, 1, -1, D), vb.view(B, 1, -1, D),
                               bool(is_causal))
        o = o.view(B, Nq, D)
        L = L.view(B, Nq)
        ctx.save_for_backward(qb, kb, vb, o, L)
        ctx.is_causal = bool(is_causal)
        ctx.in_dtype = in_dtype
        ctx.shapes = (q.shape, k.shape, v.shape)
        out = o.to(in_dtype)
        return out[0] if squeeze else out.reshape(q.shape[:-1] + (D,))

    @staticmethod
    def backward(ctx, do: Tensor):
        from . import kernels

        qb, kb, vb, o, L = ctx.saved_tensors
        B, Nq, D = qb.shape
        do3, _ = _as3d(do)
        dob = do3.to(torch.bfloat16).contiguous()
        dq, dk, dv = kernels.fa2_bwd(qb.view(B, 1, Nq, D), kb.view(B, 1, -1, D), vb.view(B, 1, -1, D),
                                     o.view(B, 1, Nq, D), L.view(B, 1, Nq), dob.view(B, 1, Nq, D),
                                     ctx.is_causal)
        qs, ks, vs = ctx.shapes
        dt = ctx.in_dtype
        return (dq.view(B, Nq, D).to(dt).reshape(qs), dk.view(B, -1, D).to(dt).reshape(ks),
                dv.view(B, -1, D).to(dt).reshape(vs), None)


def flash_attention(q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False) -> Tensor:
    """Dispatch: sm_100a kernels for CUDA tensors, the tiled PyTorch Function otherwise."""
    if _native.use_native(q, k, v):
        return FlashAttentionSM100.apply(q, k, v, is_causal)
    return FlashAttentionTorch.apply(q, k, v, is_causal)
