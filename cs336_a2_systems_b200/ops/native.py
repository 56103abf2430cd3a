"""Autograd wrappers around the sm_100a kernels, and the bf16 weight-shadow registry.

Design notes
------------
* **Weight shadows.**  Parameters stay fp32 (master weights, exactly like the reference under
  ``torch.autocast``), the tcgen05 GEMMs read bf16 copies.  A shadow is cached per *group* of
  weights that are applied to the same input (q/k/v, w1/w3 → one concatenated ``[ΣN, K]``
  buffer → ONE GEMM) and is refreshed only when a member's ``_version`` changed; the fused
  AdamW kernel writes the shadows itself, so in steady state no cast kernel runs at all.
* **Gradient destinations.**  Weight gradients are produced straight into the slot registered in
  ``grad_slots`` (a DDP bucket in NVLink symmetric memory or a flat optimizer arena) by the wgrad
  GEMM epilogue; autograd adopts the returned view as ``.grad`` without copying.
* Every Function here has a plain-PyTorch oracle in ``functional.py``; ``tests/test_kernels_gpu.py``
  compares them.
"""

from __future__ import annotations

import torch
from torch import Tensor

from . import grad_slots, kernels

BF16 = torch.bfloat16


# ------------------------------------------------------------------------------------------------
# weight shadows
# ------------------------------------------------------------------------------------------------
class _ShadowEntry:
    __slots__ = ("tensor", "versions", "row_offsets")

    def __init__(self, tensor, versions, row_offsets):
        self.tensor, self.versions, self.row_offsets = tensor, versions, row_offsets


_shadow_provider = None


def set_shadow_provider(fn) -> None:
    """``fn(weights) -> bf16 [ΣN, K] tensor | None``: lets an owner of special memory (the NVLink
    symmetric shadow arena of the fused sharded optimizer) supply the storage of new shadows."""
    global _shadow_provider
    _shadow_provider = fn


def drop_shadows(params) -> None:
    """Forget cached shadows of ``params`` (they are rebuilt, through the provider, on next use)."""
    for w in params:
        for attr in ("_b200_shadows", "_b200_shadow_slices"):
            if hasattr(w, attr):
                delattr(w, attr)


def _shadow_entry(weights: tuple[Tensor, ...]) -> _ShadowEntry:
    key = tuple(id(w) for w in weights)
    w0 = weights[0]
    reg = getattr(w0, "_b200_shadows", None)
    if reg is None:
        reg = {}
        w0._b200_shadows = reg
    ent = reg.get(key)
    if ent is None:
        K = w0.shape[1]
        offs, n = [], 0
        for w in weights:
            assert w.dim() == 2 and w.shape[1] == K and w.dtype == torch.float32
            offs.append(n)
            n += w.shape[0]
        buf = _shadow_provider(weights) if _shadow_provider is not None else None
        if buf is None:
            buf = torch.empty(n, K, dtype=BF16, device=w0.device)
        ent = _ShadowEntry(buf, [None] * len(weights), offs)
        reg[key] = ent
        for i, w in enumerate(weights):  # let the optimizer find the slice it must keep fresh
            lst = getattr(w, "_b200_shadow_slices", None)
            if lst is None:
                lst = []
                w._b200_shadow_slices = lst
            lst.append((ent, i))
    return ent


def weight_shadow(weights: tuple[Tensor, ...]) -> Tensor:
    """bf16 ``[ΣN, K]`` concatenation of the group's weights, refreshed if any member changed."""
    ent = _shadow_entry(weights)
    for i, w in enumerate(weights):
        if ent.versions[i] != w._version or ent.tensor.device != w.device:
            if ent.tensor.device != w.device:
                ent.tensor = torch.empty_like(ent.tensor, device=w.device)
                ent.versions = [None] * len(weights)
            r0 = ent.row_offsets[i]
            dst = ent.tensor[r0:r0 + w.shape[0]]
            src = w.detach()
            if src.is_contiguous():
                kernels.cast_bf16(src, dst)
            else:
                dst.copy_(src)
            ent.versions[i] = w._version
    return ent.tensor


def shadow_slices_of(w: Tensor) -> list[Tensor]:
    """bf16 views that mirror parameter ``w`` (for the fused optimizer to refresh in its own pass)."""
    out = []
    for ent, i in getattr(w, "_b200_shadow_slices", []):
        if ent.tensor.device == w.device and ent.versions[i] == w._version:
            r0 = ent.row_offsets[i]
            out.append(ent.tensor[r0:r0 + w.shape[0]])
    return out


# ------------------------------------------------------------------------------------------------
# linear
# ------------------------------------------------------------------------------------------------
def _to_bf16_2d(x: Tensor, K: int) -> Tensor:
    x2 = x.reshape(-1, K)
    if x2.dtype == torch.float32:
        return kernels.cast_bf16(x2.contiguous())
    if x2.dtype != BF16:
        x2 = x2.to(BF16)
    if x2.stride(1) != 1 or x2.stride(0) % 8 != 0 or x2.data_ptr() % 16 != 0:
        x2 = x2.contiguous()
    return x2


class LinearFn(torch.autograd.Function):
    """``y = x · [W_0; W_1; …]ᵀ`` on the tcgen05 GEMM; backward = dgrad (NN) + wgrad (TN).

    The wgrad of a fused group is ONE GEMM when the members' gradient slots are adjacent (flat arena in
    parameter order), else one per weight.  ``rope`` fuses RoPE into the forward epilogue."""

    @staticmethod
    def forward(ctx, x: Tensor, wb: Tensor, rope, *weights: Tensor):
        N, K = wb.shape
        x2 = _to_bf16_2d(x, K)
        y = kernels.gemm(x2, wb, rope=rope)
        ctx.save_for_backward(x2, wb)
        ctx.weights = weights
        ctx.x_shape = x.shape
        ctx.x_dtype = x.dtype
        return y.view(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, dy: Tensor):
        x2, wb = ctx.saved_tensors
        N, K = wb.shape
        dy2 = _to_bf16_2d(dy, N)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = kernels.gemm(dy2, wb, b_mn=True)  # [T,N]·[N,K]
            dx = dx.view(ctx.x_shape)
            if ctx.x_dtype != BF16:
                dx = dx.to(ctx.x_dtype)
        return (dx, None, None, *_weight_grads(ctx.weights, ctx.needs_input_grad[3:], dy2, x2, wb))


def _weight_grads(ws, need, dy2: Tensor, x2: Tensor, wb: Tensor) -> list:
    """Weight gradients of a (fused group of) linear map(s) ``y = x2 · wbᵀ`` for ``dy2``: one entry per weight.
    Either the optimizer runs inside the wgrad GEMM (all entries None), or the group's slots are adjacent (one GEMM),
    or one wgrad GEMM per weight straight into its gradient slot."""
    # optimizer in the wgrad epilogue (AdamW.fuse_into_backward): the whole group must belong to one fused group,
    # each weight mirrored by exactly this shadow, nothing accumulated into .grad yet
    grp = getattr(ws[0], "_b200_bwd_opt", None)
    if grp is not None and all(need) and all(getattr(w, "_b200_bwd_opt", None) is grp and w.grad is None and
                                             len(getattr(w, "_b200_shadow_slices", ())) == 1 for w in ws) and \
            wb.is_contiguous() and not torch.is_grad_enabled():
        grp.wgrad(dy2, x2, ws, wb)
        return [None] * len(ws)
    if len(ws) > 1 and all(need):
        joint = grad_slots.claim_group(ws)
        if joint is not None:
            kernels.wgrad(dy2, x2, out=joint)
            dws, r0 = [], 0
            for w in ws:
                dws.append(joint[r0:r0 + w.shape[0]])
                r0 += w.shape[0]
            return dws
    dws = []
    r0 = 0
    for i, w in enumerate(ws):
        n_i = w.shape[0]
        if need[i]:
            dst = grad_slots.claim(w)
            dyi = dy2[:, r0:r0 + n_i]  # column slice: MN-major A operand with lda = N
            if dyi.data_ptr() % 16 != 0:
                dyi = dyi.contiguous()
            dws.append(kernels.wgrad(dyi, x2, out=dst))
        else:
            dws.append(None)
        r0 += n_i
    return dws


def linear(x: Tensor, weights: tuple[Tensor, ...], rope: tuple | None = None) -> Tensor:
    """Apply one or several bias-free linear maps that share the input (outputs concatenated).
    ``rope=(cache, S, head_dim, n_cols)`` rotates the first ``n_cols`` output columns (fused QKV → q,k)."""
    wb = weight_shadow(weights)
    return LinearFn.apply(x, wb, rope, *weights)


# ------------------------------------------------------------------------------------------------
# RMSNorm (+ residual)
# ------------------------------------------------------------------------------------------------
class RMSNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor, w: Tensor, eps: float, out_dtype: torch.dtype):
        xc = x.contiguous()
        y, rstd, _ = kernels.rmsnorm_fwd(xc, w.detach(), eps, out_dtype)
        ctx.save_for_backward(xc, w, rstd)
        return y

    @staticmethod
    def backward(ctx, dy: Tensor):
        x, w, rstd = ctx.saved_tensors
        xf = x if x.dtype == torch.float32 else x.float()
        dst = grad_slots.claim(w) if ctx.needs_input_grad[1] else None
        dx, _, dw = kernels.rmsnorm_bwd(dy.contiguous(), xf, w.detach(), rstd, None, False, dw_out=dst)
        if x.dtype != torch.float32:
            dx = dx.to(x.dtype)
        return dx, (dw if ctx.needs_input_grad[1] else None), None, None


def rmsnorm(x: Tensor, w: Tensor, eps: float, out_dtype: torch.dtype) -> Tensor:
    return RMSNormFn.apply(x, w, eps, out_dtype)


class AddRMSNormFn(torch.autograd.Function):
    """``x' = x + y`` (fp32 residual stream + bf16 branch), ``h = rmsnorm(x')·w`` in bf16 — one kernel.
    Backward merges the residual-path gradient with the norm gradient and emits both the fp32
    stream gradient and its bf16 copy for the branch in one pass."""

    @staticmethod
    def forward(ctx, x: Tensor, y: Tensor, w: Tensor, eps: float):
        ctx.set_materialize_grads(False)
        h, rstd, xout = kernels.rmsnorm_fwd(x.contiguous(), w.detach(), eps, BF16, resid=y.contiguous())
        ctx.save_for_backward(xout, w, rstd)
        return xout, h

    @staticmethod
    def backward(ctx, dxout: Tensor | None, dh: Tensor | None):
        xout, w, rstd = ctx.saved_tensors
        if dh is None:
            if dxout is None:
                return None, None, None, None
            return dxout, dxout.to(BF16), None, None
        dst = grad_slots.claim(w) if ctx.needs_input_grad[2] else None
        dres = dxout.contiguous() if dxout is not None else None
        dx, dxb, dw = kernels.rmsnorm_bwd(dh.contiguous(), xout, w.detach(), rstd, dres, True, dw_out=dst)
        return dx, dxb, (dw if ctx.needs_input_grad[2] else None), None


def add_rmsnorm(x: Tensor, y: Tensor, w: Tensor, eps: float):
    return AddRMSNormFn.apply(x, y, w, eps)


class ResidualAddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x: Tensor, y: Tensor):
        return kernels.residual_add(x.contiguous(), y.contiguous())

    @staticmethod
    def backward(ctx, g: Tensor):
        return g, g.to(BF16)


def residual_add(x: Tensor, y: Tensor) -> Tensor:
    return ResidualAddFn.apply(x, y)


# ------------------------------------------------------------------------------------------------
# attention core: RoPE (in place) + FlashAttention-2
# ------------------------------------------------------------------------------------------------
class AttentionCoreFn(torch.autograd.Function):
    """``qkv``: bf16 ``[B, S, 3, H, dh]`` straight from the fused QKV GEMM.  Rotates q,k in place
    (nothing else reads the un-rotated values), runs causal FA2 on strided views (no head
    transposes), returns ``[B, S, H·dh]``.  Backward writes dq/dk/dv into one ``[B,S,3,H,dh]``
    buffer and un-rotates dq,dk in place — ready to be the dgrad/wgrad GEMM operand."""

    @staticmethod
    def forward(ctx, qkv: Tensor, rope_cache: Tensor, rope_done: bool = False):
        B, S, _, H, dh = qkv.shape
        assert qkv.is_contiguous()
        if not rope_done:  # otherwise the QKV GEMM epilogue already rotated q and k
            kernels.rope_qk_inplace(qkv, rope_cache)
        q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))  # (B, H, S, dh) strided views
        o = torch.empty(B, S, H, dh, dtype=BF16, device=qkv.device)
        _, lse = kernels.fa2_fwd(q, k, v, True, out=o.permute(0, 2, 1, 3))
        ctx.save_for_backward(qkv, o, lse, rope_cache)
        return o.view(B, S, H * dh)

    @staticmethod
    def backward(ctx, do: Tensor):
        qkv, o, lse, rope_cache = ctx.saved_tensors
        B, S, _, H, dh = qkv.shape
        q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        dqkv = torch.empty_like(qkv)
        dq, dk, dv = (dqkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
        do4 = do.contiguous().view(B, S, H, dh).permute(0, 2, 1, 3)
        # dq, dk are un-rotated inside the backward kernels (dq in its fp32→bf16 pass, dk in the epilogue)
        kernels.fa2_bwd(q, k, v, o.permute(0, 2, 1, 3), lse, do4, True, dq=dq, dk=dk, dv=dv, rope_cache=rope_cache)
        return dqkv, None, None


def attention_core(qkv: Tensor, rope_cache: Tensor, rope_done: bool = False) -> Tensor:
    return AttentionCoreFn.apply(qkv, rope_cache, rope_done)


# ------------------------------------------------------------------------------------------------
# SwiGLU gate, embedding, cross-entropy
# ------------------------------------------------------------------------------------------------
class SwiGLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ab: Tensor):
        abc = ab.contiguous()
        ctx.save_for_backward(abc)
        return kernels.swiglu_fwd(abc)

    @staticmethod
    def backward(ctx, dg: Tensor):
        (ab,) = ctx.saved_tensors
        return kernels.swiglu_bwd(ab, dg)


def swiglu(ab: Tensor) -> Tensor:
    return SwiGLUFn.apply(ab)


class SwiGLUFFNFn(torch.autograd.Function):
    """The whole feed-forward block ``w2(silu(w1 x) * w3 x)`` with the gate inside the GEMM epilogues:

    forward   ``g, ab = gemm_swiglu_fwd(x, [w1; w3])`` — ONE GEMM writes the gated product and the pre-activations;
              ``y = g · w2ᵀ``
    backward  ``dab = gemm_swiglu_bwd(dy, w2, ab)`` — the dgrad GEMM of w2 applies the gate's backward in its epilogue
              (``dg`` never reaches memory); ``dx = dab · [w1; w3]``; wgrads of w2 (from ``g``) and w1‖w3 (from ``dab``)

    Saved for backward: ``x``, ``ab``, ``g`` and the two weight shadows — what the unfused chain saves too."""

    #: apply the gate's backward inside the w2 dgrad GEMM (True) or as the separate elementwise kernel (False).
    #: Measured on B200 (profiles/r2): the forward fusion saves ~7 us per layer; the backward epilogue (two dependent
    #: 16-byte loads per 8 elements) makes the dgrad GEMM epilogue-bound and loses more than the saved kernel costs.
    FUSE_BACKWARD = False

    @staticmethod
    def forward(ctx, x: Tensor, wb13: Tensor, wb2: Tensor, w1: Tensor, w3: Tensor, w2: Tensor):
        K = wb13.shape[1]
        x2 = _to_bf16_2d(x, K)
        g, ab = kernels.gemm_swiglu_fwd(x2, wb13)
        y = kernels.gemm(g, wb2)
        ctx.save_for_backward(x2, ab, g, wb13, wb2)
        ctx.weights = (w1, w3, w2)
        ctx.x_shape, ctx.x_dtype = x.shape, x.dtype
        return y.view(*x.shape[:-1], wb2.shape[0])

    @staticmethod
    def backward(ctx, dy: Tensor):
        x2, ab, g, wb13, wb2 = ctx.saved_tensors
        w1, w3, w2 = ctx.weights
        dy2 = _to_bf16_2d(dy, wb2.shape[0])
        need = ctx.needs_input_grad
        if SwiGLUFFNFn.FUSE_BACKWARD:
            dab = kernels.gemm_swiglu_bwd(dy2, wb2, ab)
        else:
            dab = kernels.swiglu_bwd(ab, kernels.gemm(dy2, wb2, b_mn=True))
        dx = None
        if need[0]:
            dx = kernels.gemm(dab, wb13, b_mn=True).view(ctx.x_shape)
            if ctx.x_dtype != BF16:
                dx = dx.to(ctx.x_dtype)
        (dw2,) = _weight_grads((w2,), (need[5],), dy2, g, wb2)
        dw1, dw3 = _weight_grads((w1, w3), (need[3], need[4]), dab, x2, wb13)
        return dx, None, None, dw1, dw3, dw2


def swiglu_ffn(x: Tensor, w1: Tensor, w3: Tensor, w2: Tensor) -> Tensor | None:
    """Fused training path of the feed-forward block, or None when the shape cannot take it (fewer than 256 tokens,
    d_ff not a multiple of 128) — the caller then runs the unfused chain."""
    wb13 = weight_shadow((w1, w3))
    K = w1.shape[1]
    T = x.numel() // K
    if T < 256 or (2 * w1.shape[0]) % 256 != 0 or w1.shape[0] != w3.shape[0]:
        return None
    return SwiGLUFFNFn.apply(x, wb13, weight_shadow((w2,)), w1, w3, w2)


def gated_shadow(w1: Tensor, w3: Tensor) -> Tensor:
    """bf16 ``[2*d_ff, d]`` with the rows of w1 and w3 interleaved (w1[0], w3[0], w1[1], ...): the B operand of the
    GEMM whose epilogue applies ``silu(gate) * up`` (``EPI_BF16_SWIGLU``).  Derived from the always-fresh
    concatenated shadow; rebuilt only when the parameters may have changed."""
    key = (kernels.param_epoch, w1._version, w3._version, w1.data_ptr(), w3.data_ptr())
    ent = getattr(w1, "_b200_gate_shadow", None)
    if ent is None or ent[0] != key:
        ws = weight_shadow((w1, w3))  # [2F, K]: rows 0..F-1 = w1, F..2F-1 = w3
        F_ = w1.shape[0]
        buf = ent[1] if ent is not None and ent[1].shape == ws.shape and ent[1].device == ws.device \
            else torch.empty_like(ws)
        buf.view(F_, 2, -1).copy_(ws.view(2, F_, -1).transpose(0, 1))
        ent = (key, buf)
        w1._b200_gate_shadow = ent
    return ent[1]


def swiglu_linear_nograd(x: Tensor, w1: Tensor, w3: Tensor) -> Tensor:
    """``silu(x w1ᵀ) * (x w3ᵀ)`` as ONE GEMM with the gate in its epilogue.  Nothing is saved for
    backward, so this is the serving / evaluation path (``torch.no_grad()``); training keeps the two-kernel form
    because its backward needs both pre-activations."""
    K = w1.shape[1]
    x2 = _to_bf16_2d(x, K)
    h = kernels.gemm(x2, gated_shadow(w1, w3), epilogue=kernels.EPI_BF16_SWIGLU)
    return h.view(*x.shape[:-1], w1.shape[0])


class EmbeddingFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ids: Tensor, w: Tensor):
        ctx.save_for_backward(ids)
        ctx.w = w
        return kernels.embedding_fwd(ids, w.detach())

    @staticmethod
    def backward(ctx, dout: Tensor):
        (ids,) = ctx.saved_tensors
        w = ctx.w
        dst = grad_slots.claim(w)
        if dst is None:
            dst = torch.empty_like(w)
        kernels.embedding_bwd(ids, dout.float() if dout.dtype != torch.float32 else dout, dst, zero_first=True)
        return None, dst


def embedding(ids: Tensor, w: Tensor) -> Tensor:
    return EmbeddingFn.apply(ids, w)


class CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits: Tensor, targets: Tensor):
        V = logits.shape[-1]
        l2 = logits.reshape(-1, V)
        if l2.stride(1) != 1:
            l2 = l2.contiguous()
        t = targets.reshape(-1).contiguous()
        loss, lse = kernels.xent_fwd(l2, t)
        ctx.save_for_backward(l2, t, lse)
        ctx.shape = logits.shape
        return loss

    @staticmethod
    def backward(ctx, dloss: Tensor):
        l2, t, lse = ctx.saved_tensors
        return kernels.xent_bwd(l2, t, lse, dloss).view(ctx.shape), None


def cross_entropy(logits: Tensor, targets: Tensor) -> Tensor:
    return CrossEntropyFn.apply(logits, targets)


# ------------------------------------------------------------------------------------------------
# row softmax, materialised attention (comparison arm of the attention benchmarks)
# ------------------------------------------------------------------------------------------------
class SoftmaxFn(torch.autograd.Function):
    """Softmax over the last dimension on the row kernel (``csrc/softmax.cu``)."""

    @staticmethod
    def forward(ctx, x: Tensor):
        x2 = x.reshape(-1, x.shape[-1])
        if x2.stride(1) != 1:
            x2 = x2.contiguous()
        y = kernels.softmax_rows(x2)
        ctx.save_for_backward(y)
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy: Tensor):
        (y,) = ctx.saved_tensors
        dy2 = dy.reshape(y.shape).to(y.dtype).contiguous()
        return kernels.softmax_bwd_rows(y, dy2).view(dy.shape)


def softmax_lastdim(x: Tensor) -> Tensor:
    return SoftmaxFn.apply(x)


class MaterializedAttentionFn(torch.autograd.Function):
    """Attention that materialises the ``N_q x N_k`` probabilities, on this library's kernels: per batch element
    ``S = Q Kᵀ`` (tcgen05 GEMM), one fused scale + causal-mask + softmax row kernel, ``O = P V`` (GEMM); backward =
    four GEMMs and the softmax-backward row kernel.  The counterpart of what the reference's eager / ``torch.compile``d
    ``scaled_dot_product_attention`` executes (``model.py:400-432``, ``benchmark_attention.py:113-201``); it is the
    O(N²)-memory arm FlashAttention is compared against, not a path the LM uses."""

    @staticmethod
    def forward(ctx, q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False):
        B, Nq, D = q.shape
        Nk = k.shape[1]
        if q.dtype == torch.float32:
            qb, kb, vb = kernels.cast3_bf16(q.contiguous(), k.contiguous(), v.contiguous())
        else:
            qb, kb, vb = q.contiguous(), k.contiguous(), v.contiguous()
        scale = 1.0 / (D ** 0.5)
        P = torch.empty(B, Nq, Nk, dtype=BF16, device=q.device)
        O = torch.empty(B, Nq, D, dtype=BF16, device=q.device)
        for b in range(B):
            kernels.gemm(qb[b], kb[b], out=P[b])                                     # S (bf16) into P's storage
            kernels.softmax_rows(P[b], scale=scale, causal_period=Nq if is_causal else 0, out=P[b])
            kernels.gemm(P[b], vb[b], b_mn=True, out=O[b])
        ctx.save_for_backward(qb, kb, vb, P)
        ctx.scale = scale
        ctx.in_dtype = q.dtype
        return O.to(q.dtype) if q.dtype != BF16 else O

    @staticmethod
    def backward(ctx, dO: Tensor):
        qb, kb, vb, P = ctx.saved_tensors
        B, Nq, D = qb.shape
        dOb = dO.to(BF16).contiguous()
        dq, dk, dv = torch.empty_like(qb), torch.empty_like(kb), torch.empty_like(vb)
        dP = torch.empty_like(P[0])
        for b in range(B):
            kernels.gemm(P[b], dOb[b], a_mn=True, b_mn=True, out=dv[b])             # dV = Pᵀ dO
            kernels.gemm(dOb[b], vb[b], out=dP)                                      # dP = dO Vᵀ
            kernels.softmax_bwd_rows(P[b], dP, scale=ctx.scale, out=dP)              # dS (scaled), in place
            kernels.gemm(dP, kb[b], b_mn=True, out=dq[b])                            # dQ = dS K
            kernels.gemm(dP, qb[b], a_mn=True, b_mn=True, out=dk[b])                 # dK = dSᵀ Q
        dt = ctx.in_dtype
        return dq.to(dt), dk.to(dt), dv.to(dt), None


def materialized_attention(q: Tensor, k: Tensor, v: Tensor, is_causal: bool = False) -> Tensor:
    return MaterializedAttentionFn.apply(q, k, v, is_causal)


# ------------------------------------------------------------------------------------------------
# optimizer-side helpers
# ------------------------------------------------------------------------------------------------
def clip_grad_norm_(grads: list[Tensor], max_norm: float, eps: float = 1e-6) -> Tensor:
    assert all(g.is_contiguous() for g in grads), "native clip needs contiguous gradients (api.clip_gradient checks)"
    return kernels.clip_grad_norm(list(grads), max_norm, eps)


class AdamWPlan:
    """One fused launch over a fixed set of (param, grad, m, v) tensors (+ their bf16 shadows)."""

    def __init__(self, params, grads, ms, vs, key, step0: int, mode: int = 0):
        self.key = key
        self.params = list(params)
        self.host_step = step0  # the host's idea of the device step counter (see AdamW._step_fused)
        flat_p, flat_g, flat_m, flat_v, flat_s = [], [], [], [], []
        self._extra_shadows = []
        for p_, g_, m_, v_ in zip(params, grads, ms, vs):
            sl = shadow_slices_of(p_)
            flat_p.append(p_.data)
            flat_g.append(g_)
            flat_m.append(m_)
            flat_v.append(v_)
            first = sl[0] if sl else None
            if first is not None and not first.is_contiguous():
                first = None
            flat_s.append(first)
            for extra in sl[1:]:
                self._extra_shadows.append((p_, extra))
        self._launch = kernels.AdamWLaunch(flat_p, flat_g, flat_m, flat_v, flat_s, step0, mode)

    def set_hyper(self, lr, b1, b2, eps, wd, grad_scale: float = 1.0) -> None:
        self._launch.set_hyper(lr, b1, b2, eps, wd, grad_scale)

    def launch(self, max_blocks: int = 0) -> None:
        self._launch.launch(max_blocks)
        for p_, sh in self._extra_shadows:  # a weight mirrored in several groups (rare)
            kernels.cast_bf16(p_.data, sh)

    def device_step(self) -> int:
        return int(self._launch.step.item())
