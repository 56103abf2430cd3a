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
    """FA2 on the tcgen05 kernels.  ``(B, N, D)`` inputs, bf16 tensor-core operands, fp32 softmax/accumulate.

    bf16 inputs go straight to the kernels.  fp32 inputs (the reference's tests feed fp32,
    ``tests/test_attention.py:29-40``) cost ONE extra launch: a fused cast of q, k, v to the bf16 operands the
    tensor cores take; the forward kernel then writes ``O`` in fp32 directly, the backward's pre-pass reads the fp32
    ``O`` / ``dO`` as they are, and dQ/dK/dV come out of the kernels in fp32 — no ``Tensor.to`` anywhere.
    ``L`` is always fp32.  Saved tensors: ``Q, K, V`` (bf16 operands), ``O`` (caller's dtype), ``L``.
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
        f32 = in_dtype == torch.float32 and k.dtype == torch.float32 and v.dtype == torch.float32
        if f32:
            qb, kb, vb = kernels.cast3_bf16(q3.contiguous(), k3.contiguous(), v3.contiguous())
        else:
            qb, kb, vb = (t if t.dtype == torch.bfloat16 else t.to(torch.bfloat16) for t in (q3, k3, v3))
            qb, kb, vb = (t if t.stride(-1) == 1 else t.contiguous() for t in (qb, kb, vb))
        B, Nq, D = qb.shape

        def bhnd(t):  # kernels take (B, H, N, D) views with arbitrary batch/row strides; here H = 1
            return t.unsqueeze(1)

        o, L = kernels.fa2_fwd(bhnd(qb), bhnd(kb), bhnd(vb), bool(is_causal), out_f32=f32)
        o = o.squeeze(1)
        L = L.view(B, Nq)
        ctx.save_for_backward(qb, kb, vb, o, L)
        ctx.is_causal = bool(is_causal)
        ctx.in_dtype = in_dtype
        ctx.f32 = f32
        ctx.shapes = (q.shape, k.shape, v.shape)
        out = o if o.dtype == in_dtype else o.to(in_dtype)
        return out[0] if squeeze else out.reshape(q.shape[:-1] + (D,))

    @staticmethod
    def backward(ctx, do: Tensor):
        from . import kernels

        qb, kb, vb, o, L = ctx.saved_tensors
        B, Nq, D = qb.shape
        do3, _ = _as3d(do)
        if do3.dtype not in (torch.bfloat16, torch.float32):
            do3 = do3.to(torch.bfloat16)
        if do3.stride(-1) != 1:
            do3 = do3.contiguous()
        dq, dk, dv = kernels.fa2_bwd(qb.unsqueeze(1), kb.unsqueeze(1), vb.unsqueeze(1), o.unsqueeze(1),
                                     L.view(B, 1, Nq), do3.unsqueeze(1), ctx.is_causal, grads_f32=ctx.f32)
        qs, ks, vs = ctx.shapes
        dt = ctx.in_dtype
        fix = lambda g, shape: (g.squeeze(1) if g.dtype == dt else g.squeeze(1).to(dt)).reshape(shape)
        return fix(dq, qs), fix(dk, ks), fix(dv, vs), None


def flash_attention(q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False) -> Tensor:
    """Dispatch: sm_100a kernels for CUDA tensors, the tiled PyTorch Function otherwise."""
    if _native.use_native(q, k, v):
        return FlashAttentionSM100.apply(q, k, v, is_causal)
    return FlashAttentionTorch.apply(q, k, v, is_causal)
