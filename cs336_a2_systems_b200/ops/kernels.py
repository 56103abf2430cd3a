"""ctypes bindings of ``libb200sys.so`` (see ``include/b200sys.h``).

Thin, allocation-aware wrappers: they validate layouts, allocate outputs / scratch with torch
(so the caching allocator and CUDA-graph private pools apply), pass raw device pointers and the
current stream, and raise on any non-zero return code.  No autograd here — see ``native.py``.
"""

from __future__ import annotations

import ctypes
from os import environ as _environ
from ctypes import c_float, c_int, c_longlong, c_void_p

import torch
from torch import Tensor

from .. import _native
from .._native import check, p, stream_ptr

BF16, F32 = 0, 1
_DT = {torch.bfloat16: BF16, torch.float32: F32}

EPI_BF16, EPI_F32, EPI_F32_ACC, EPI_BF16_SWIGLU, EPI_F32_RESID, EPI_BF16_ROPE, EPI_F32_T, EPI_ADAMW = 0, 1, 2, 3, 4, 5, 6, 7
EPI_SWIGLU_FWD, EPI_SWIGLU_BWD = 8, 9

#: route eligible GEMMs through the CTA-pair (cta_group::2) kernel
GEMM_PAIR = _environ.get("B200SYS_GEMM_PAIR", "auto")  # "auto": per-shape choice, "1": always when legal, "0": never

#: launch counter (kernels launched through this module), for bench.py's "gpu_launches"
launches = 0


#: bumped whenever a fused optimizer kernel may have rewritten parameters behind autograd's back (raw-pointer
#: launches do not touch ``Tensor._version``); caches derived from parameter values key on it
param_epoch = 0


def bump_param_epoch() -> None:
    global param_epoch
    param_epoch += 1


def native_launches() -> int:
    """Exact number of kernels the native library has launched in this process (counted at the launch sites)."""
    fn = _lib().b200_launch_count
    fn.restype = ctypes.c_longlong
    return int(fn())


def _count(n: int = 1) -> None:
    global launches
    launches += n


def _lib():
    return _native.lib()


# ------------------------------------------------------------------------------------------------
# GEMM
# ------------------------------------------------------------------------------------------------
import functools as _functools


def _load_tuned() -> dict:
    """Measured plans: ``ops/gemm_tuned.json`` maps "M,N,K,a_mn,b_mn" to [use_pair, tile width], written by
    ``python -m cs336_a2_systems_b200.bench.gemm --tune`` on a B200 (every legal kernel x tile width timed per shape).
    Shapes that are not in the table fall back to the wave model below."""
    import json as _json
    import os as _os
    try:
        with open(_os.path.join(_os.path.dirname(__file__), "gemm_tuned.json")) as f:
            return {k: (bool(v[0]), int(v[1])) for k, v in _json.load(f).get("plans", {}).items()}
    except Exception:
        return {}


_TUNED = _load_tuned() if _environ.get("B200SYS_GEMM_TUNED", "1") == "1" else {}


def plan_gemm(M: int, N: int, K: int, a_mn: bool, b_mn: bool, pair_ok: bool):
    """(use_pair, tile width) for C[M, N] = A·Bᵀ: the measured plan for this exact shape if there is one, else the model."""
    t = _TUNED.get(f"{M},{N},{K},{int(a_mn)},{int(b_mn)}")
    if t is not None and (pair_ok or not t[0]):
        return t
    return _plan_gemm(M, N, b_mn, pair_ok)


@_functools.lru_cache(maxsize=None)
def _plan_gemm(M: int, N: int, b_mn: bool, pair_ok: bool, sms: int = 148):
    """Pick (use_pair, tile width) by a wave model calibrated on B200 (profiles/gemm_*): a 1-CTA tile of width BN
    costs BN·{1.0, 1.05, 1.38} for BN = 256/192/128 (narrow tiles are issue-bound); a CTA-pair tile (256 rows)
    costs BN·0.9 on half as many schedulable units."""
    ceil = lambda a, b: -(-a // b)
    nm = ceil(M, 128)
    best = (False, 256, float("inf"))
    for bn, pen in ((256, 1.0), (192, 1.05), (128, 1.38)):
        cost = ceil(nm * ceil(N, bn), sms) * bn * pen
        if cost < best[2]:
            best = (False, bn, cost)
    if pair_ok:
        for bn, pen in ((256, 0.90), (192, 0.93), (128, 1.05)):
            if b_mn and bn % 128:
                continue
            cost = ceil(ceil(nm, 2) * ceil(N, bn), sms // 2) * bn * pen
            if cost < best[2]:
                best = (True, bn, cost)
    return best[0], best[1]


def gemm_supported(a: Tensor, b: Tensor) -> bool:
    """TMA needs 16-byte aligned bases and row strides, unit inner stride."""
    for t in (a, b):
        if t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
            return False
        if t.stride(0) % 8 != 0 or t.data_ptr() % 16 != 0:
            return False
    return True


def wgrad(dy: Tensor, x: Tensor, out: Tensor | None = None) -> Tensor:
    """``dW[N_out, K_in] = dyᵀ · x`` (fp32).  Always computed as the transpose ``xᵀ · dy`` with the transposed-store
    epilogue (``EPI_F32_T``): a thread owns one row of dWᵀ, so for every accumulator register the 32 lanes of a warp
    write 32 consecutive floats of one dW row — one coalesced 128-byte store, where the row-per-thread fp32 epilogue
    scatters each 16-byte vector store over 32 cache lines (measured 0.72-0.80x of cuBLAS on the square / wide
    wgrads, `profiles/r2/gemm_vs_cublas_xl.json`)."""
    n_out, k_in = dy.shape[1], x.shape[1]
    if out is None:
        out = torch.empty(n_out, k_in, dtype=torch.float32, device=dy.device)
    if os_environ_flag("B200SYS_WGRAD_PLAIN"):
        return gemm(dy, x, a_mn=True, b_mn=True, epilogue=EPI_F32, out=out)
    return gemm(x, dy, a_mn=True, b_mn=True, epilogue=EPI_F32_T, out=out)


def os_environ_flag(name: str) -> bool:
    return _environ.get(name, "0") == "1"


def gemm(a: Tensor, b: Tensor, *, a_mn: bool = False, b_mn: bool = False, epilogue: int = EPI_BF16,
         out: Tensor | None = None, aux: Tensor | None = None, force_bn: int = 0,
         rope: tuple | None = None, pair: bool | None = None) -> Tensor:
    """``C = A · Bᵀ``.  ``a``: [M,K] (or [K,M] when ``a_mn``); ``b``: [N,K] (or [K,N] when ``b_mn``).

    ``rope=(cache, S, head_dim, n_cols)`` selects the RoPE epilogue: columns ``< n_cols`` of the bf16 output
    are rotated (row ``r`` is position ``r % S``)."""
    rope_S = rope_dh = rope_cols = rope_ctx = 0
    if rope is not None:
        cache, rope_S, rope_dh, rope_cols = rope
        assert cache.dtype == torch.float32 and cache.is_contiguous() and cache.shape[2] * 2 == rope_dh
        rope_ctx = cache.shape[1]
        aux, epilogue = cache, EPI_BF16_ROPE
    M, K = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    if K != Kb:
        raise ValueError(f"gemm: inner dimensions differ ({K} vs {Kb})")
    if not gemm_supported(a, b):
        raise ValueError("gemm: operands must be bf16, 2-D, unit inner stride, 16-byte aligned rows")
    n_out = N // 2 if epilogue == EPI_BF16_SWIGLU else N
    out_shape = (N, M) if epilogue == EPI_F32_T else (M, n_out)
    if out is None:
        dt = torch.bfloat16 if epilogue in (EPI_BF16, EPI_BF16_SWIGLU, EPI_BF16_ROPE) else torch.float32
        out = torch.empty(out_shape, dtype=dt, device=a.device)
    else:
        assert tuple(out.shape) == out_shape and out.stride(1) == 1
    ld_aux = aux.stride(0) if (aux is not None and rope is None) else 0
    # CTA-pair kernel (tcgen05 cta_group::2, 256-row tiles): plain / fp32 / RoPE / transposed epilogues
    pair_ok = epilogue in (EPI_BF16, EPI_F32, EPI_BF16_ROPE, EPI_F32_T) and not (a_mn and not b_mn) and M >= 256
    if pair is None and GEMM_PAIR == "auto" and force_bn == 0:
        use_pair, force_bn = plan_gemm(M, N, K, a_mn, b_mn, pair_ok)
    else:
        use_pair = pair_ok and (pair if pair is not None else GEMM_PAIR == "1")
    fn = _lib().b200_gemm2_bf16_ex if use_pair else _lib().b200_gemm_bf16_ex
    check(fn(p(a), p(b), p(out), p(aux), c_int(M), c_int(N), c_int(K), c_int(a.stride(0)),
                                   c_int(b.stride(0)), c_int(out.stride(0)), c_int(ld_aux), c_int(int(a_mn)),
                                   c_int(int(b_mn)), c_int(epilogue), c_int(force_bn), c_int(rope_S), c_int(rope_dh),
                                   c_int(rope_cols), c_int(rope_ctx), stream_ptr()), "gemm")
    _count()
    return out


def swiglu_up_supported(x2: Tensor, w13: Tensor) -> bool:
    """The fused up-projection needs CTA pairs (>= 256 rows) and d_ff a multiple of 128."""
    F2 = w13.shape[0]
    return x2.shape[0] >= 256 and F2 % 256 == 0 and gemm_supported(x2, w13)


def gemm_swiglu_fwd(x2: Tensor, w13: Tensor):
    """``x2``: bf16 [T, K]; ``w13``: bf16 [2F, K] = w1 stacked on w3.  ONE GEMM whose epilogue applies the gate:
    returns ``(g, ab)`` with ``g = silu(x w1ᵀ) * (x w3ᵀ)`` (bf16 [T, F]) and the pre-activations ``ab`` (bf16 [T, 2F],
    ``a | b``) the backward needs (``EPI_SWIGLU_FWD``: each 256-column tile is 128 rows of w1 next to the same 128
    rows of w3, so no interleaved weight copy is needed)."""
    T, K = x2.shape
    F2 = w13.shape[0]
    F = F2 // 2
    g = torch.empty(T, F, dtype=torch.bfloat16, device=x2.device)
    ab = torch.empty(T, F2, dtype=torch.bfloat16, device=x2.device)
    check(_lib().b200_gemm2_bf16_ex(p(x2), p(w13), p(g), p(ab), c_int(T), c_int(F2), c_int(K), c_int(x2.stride(0)),
                                    c_int(w13.stride(0)), c_int(F), c_int(F2), c_int(0), c_int(0),
                                    c_int(EPI_SWIGLU_FWD), c_int(256), c_int(0), c_int(0), c_int(0), c_int(0),
                                    stream_ptr()), "gemm_swiglu_fwd")
    _count()
    return g, ab


def gemm_swiglu_bwd(dy2: Tensor, w2: Tensor, ab: Tensor) -> Tensor:
    """``dy2``: bf16 [T, d]; ``w2``: bf16 [d, F] (the down-projection, read as an MN-major B operand); ``ab``: the saved
    pre-activations.  The dgrad GEMM ``dg = dy · w2`` with the gate's backward in its epilogue (``EPI_SWIGLU_BWD``):
    returns ``dab = (da | db)`` bf16 [T, 2F]; ``dg`` never reaches memory."""
    T, d = dy2.shape
    F = w2.shape[1]
    assert w2.shape[0] == d and ab.shape == (T, 2 * F) and ab.is_contiguous()
    dab = torch.empty(T, 2 * F, dtype=torch.bfloat16, device=dy2.device)
    pair_ok = T >= 256
    use_pair, bn = _plan_gemm(T, F, True, pair_ok) if GEMM_PAIR == "auto" else (pair_ok and GEMM_PAIR == "1", 0)
    fn = _lib().b200_gemm2_bf16_ex if use_pair else _lib().b200_gemm_bf16_ex
    check(fn(p(dy2), p(w2), p(dab), p(ab), c_int(T), c_int(F), c_int(d), c_int(dy2.stride(0)), c_int(w2.stride(0)),
             c_int(2 * F), c_int(2 * F), c_int(0), c_int(1), c_int(EPI_SWIGLU_BWD), c_int(bn), c_int(0), c_int(0),
             c_int(0), c_int(0), stream_ptr()), "gemm_swiglu_bwd")
    _count()
    return dab


# ------------------------------------------------------------------------------------------------
# RMSNorm / residual
# ------------------------------------------------------------------------------------------------
def rmsnorm_fwd(x: Tensor, w: Tensor, eps: float, out_dtype: torch.dtype, resid: Tensor | None = None,
                want_xout: bool = False):
    """Returns ``(y, rstd, xout)``; with ``resid`` (bf16) the normalised value is ``x + resid``."""
    d = x.shape[-1]
    x2 = x.reshape(-1, d)
    assert x2.is_contiguous() and w.dtype == torch.float32 and w.is_contiguous()
    T = x2.shape[0]
    y = torch.empty(x.shape, dtype=out_dtype, device=x.device)
    rstd = torch.empty(T, dtype=torch.float32, device=x.device)
    xout = torch.empty(x.shape, dtype=torch.float32, device=x.device) if (want_xout or resid is not None) else None
    if resid is not None:
        assert resid.dtype == torch.bfloat16 and resid.is_contiguous() and resid.numel() == x.numel()
    check(_lib().b200_rmsnorm_fwd(p(x2), c_int(_DT[x.dtype]), p(resid), p(w), p(y), c_int(_DT[out_dtype]), p(xout),
                                  p(rstd), c_int(T), c_int(d), c_float(eps), stream_ptr()), "rmsnorm_fwd")
    _count()
    return y, rstd, xout


def rmsnorm_bwd(dy: Tensor, x: Tensor, w: Tensor, rstd: Tensor, dres_in: Tensor | None, want_dx_bf16: bool,
                dw_out: Tensor | None = None, want_dx_f32: bool = True):
    """``x`` is the fp32 tensor that was normalised.  Returns ``(dx_f32, dx_bf16, dw)``."""
    d = x.shape[-1]
    T = x.numel() // d
    assert x.dtype == torch.float32 and x.is_contiguous() and dy.is_contiguous()
    dx = torch.empty_like(x) if want_dx_f32 else None
    dxb = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device) if want_dx_bf16 else None
    dw = dw_out if dw_out is not None else torch.empty_like(w)
    rows = _lib().b200_rmsnorm_bwd_workspace_rows(c_int(T))
    scratch = torch.empty(rows * d, dtype=torch.float32, device=x.device)
    check(_lib().b200_rmsnorm_bwd(p(dy), c_int(_DT[dy.dtype]), p(x), p(w), p(rstd), p(dres_in), p(dx), p(dxb), p(dw),
                                  c_int(0), p(scratch), c_int(T), c_int(d), stream_ptr()), "rmsnorm_bwd")
    _count(2)
    return dx, dxb, dw


def residual_add(x: Tensor, y: Tensor) -> Tensor:
    assert x.dtype == torch.float32 and y.dtype == torch.bfloat16 and x.is_contiguous() and y.is_contiguous()
    out = torch.empty_like(x)
    check(_lib().b200_residual_add(p(x), p(y), p(out), c_longlong(x.numel()), stream_ptr()), "residual_add")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# RoPE / SwiGLU / embedding / cast
# ------------------------------------------------------------------------------------------------
def rope_qk_inplace(qkv: Tensor, cache: Tensor, inverse: bool = False) -> None:
    """``qkv``: contiguous bf16 ``[B, S, 3, H, dh]``; rotates q and k in place."""
    B, S, three, H, dh = qkv.shape
    assert three == 3 and qkv.is_contiguous() and qkv.dtype == torch.bfloat16
    assert cache.dtype == torch.float32 and cache.is_contiguous() and cache.shape[2] == dh // 2
    check(_lib().b200_rope_qk_inplace(p(qkv), p(cache), c_int(cache.shape[1]), c_int(B), c_int(S), c_int(H),
                                      c_int(dh), c_int(int(inverse)), stream_ptr()), "rope")
    _count()


def swiglu_fwd(ab: Tensor) -> Tensor:
    F = ab.shape[-1] // 2
    assert ab.is_contiguous() and ab.dtype == torch.bfloat16
    g = torch.empty(ab.shape[:-1] + (F,), dtype=torch.bfloat16, device=ab.device)
    check(_lib().b200_swiglu_fwd(p(ab), p(g), c_longlong(ab.numel() // (2 * F)), c_int(F), stream_ptr()), "swiglu_fwd")
    _count()
    return g


def swiglu_bwd(ab: Tensor, dg: Tensor) -> Tensor:
    F = ab.shape[-1] // 2
    dg = dg.contiguous()
    dab = torch.empty_like(ab)
    check(_lib().b200_swiglu_bwd(p(ab), p(dg), p(dab), c_longlong(ab.numel() // (2 * F)), c_int(F), stream_ptr()),
          "swiglu_bwd")
    _count()
    return dab


def embedding_fwd(ids: Tensor, w: Tensor) -> Tensor:
    assert ids.dtype == torch.int64 and w.dtype == torch.float32 and w.is_contiguous()
    idc = ids.contiguous()
    out = torch.empty(ids.shape + (w.shape[1],), dtype=torch.float32, device=w.device)
    check(_lib().b200_embedding_fwd(p(idc), p(w), p(out), c_longlong(idc.numel()), c_int(w.shape[1]),
                                    c_int(w.shape[0]), stream_ptr()), "embedding_fwd")
    _count()
    return out


def embedding_bwd(ids: Tensor, dout: Tensor, dw: Tensor, zero_first: bool = True) -> Tensor:
    idc = ids.contiguous()
    dout = dout.contiguous()
    assert dout.dtype == torch.float32 and dw.dtype == torch.float32 and dw.is_contiguous()
    check(_lib().b200_embedding_bwd(p(idc), p(dout), p(dw), c_longlong(idc.numel()), c_int(dw.shape[1]),
                                    c_int(dw.shape[0]), c_int(int(zero_first)), stream_ptr()), "embedding_bwd")
    _count(2 if zero_first else 1)
    return dw


def cast_bf16(src: Tensor, out: Tensor | None = None) -> Tensor:
    assert src.dtype == torch.float32 and src.is_contiguous()
    if out is None:
        out = torch.empty(src.shape, dtype=torch.bfloat16, device=src.device)
    assert out.is_contiguous() and out.numel() == src.numel()
    check(_lib().b200_cast_f32_to_bf16(p(src), p(out), c_longlong(src.numel()), stream_ptr()), "cast")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# cross-entropy
# ------------------------------------------------------------------------------------------------
def xent_fwd(logits2d: Tensor, targets: Tensor):
    T, V = logits2d.shape
    assert logits2d.stride(1) == 1 and targets.dtype == torch.int64 and targets.is_contiguous()
    row_loss = torch.empty(T, dtype=torch.float32, device=logits2d.device)
    lse = torch.empty(T, dtype=torch.float32, device=logits2d.device)
    loss = torch.empty((), dtype=torch.float32, device=logits2d.device)
    check(_lib().b200_xent_fwd(p(logits2d), c_int(_DT[logits2d.dtype]), p(targets), p(row_loss), p(lse), p(loss),
                               c_int(T), c_int(V), c_int(logits2d.stride(0)), stream_ptr()), "xent_fwd")
    _count(2)
    return loss, lse


def xent_bwd(logits2d: Tensor, targets: Tensor, lse: Tensor, dloss: Tensor, out: Tensor | None = None) -> Tensor:
    T, V = logits2d.shape
    if out is None:
        out = torch.empty_like(logits2d)
    dloss = dloss.to(torch.float32).contiguous()
    check(_lib().b200_xent_bwd(p(logits2d), c_int(_DT[logits2d.dtype]), p(targets), p(lse), p(dloss), p(out),
                               c_int(T), c_int(V), c_int(logits2d.stride(0)), c_int(out.stride(0)), stream_ptr()),
          "xent_bwd")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# row softmax
# ------------------------------------------------------------------------------------------------
def softmax_rows(x2d: Tensor, out_dtype: torch.dtype | None = None, scale: float = 1.0, causal_period: int = 0,
                 out: Tensor | None = None) -> Tensor:
    """``softmax(scale * x2d)`` along the last dim of a 2-D tensor with unit inner stride; ``causal_period`` = N masks
    row ``r`` (query ``r % N``) to columns ``0..r % N``."""
    assert x2d.dim() == 2 and x2d.stride(1) == 1 and x2d.dtype in _DT
    rows, cols = x2d.shape
    out_dtype = out_dtype or x2d.dtype
    if out is None:
        out = torch.empty(rows, cols, dtype=out_dtype, device=x2d.device)
    assert out.shape == x2d.shape and out.stride(1) == 1 and out.dtype == out_dtype
    check(_lib().b200_softmax_rows(p(x2d), c_int(_DT[x2d.dtype]), p(out), c_int(_DT[out_dtype]), c_longlong(rows),
                                   c_int(cols), c_longlong(x2d.stride(0)), c_longlong(out.stride(0)), c_float(scale),
                                   c_int(causal_period), stream_ptr()), "softmax_rows")
    _count()
    return out


def softmax_bwd_rows(prob: Tensor, dprob: Tensor, scale: float = 1.0, out: Tensor | None = None) -> Tensor:
    """``scale * p * (dp - rowsum(dp * p))`` for contiguous 2-D ``p`` / ``dp`` of the same dtype (may run in place)."""
    assert prob.shape == dprob.shape and prob.dtype == dprob.dtype and prob.is_contiguous() and dprob.is_contiguous()
    rows, cols = prob.shape
    if out is None:
        out = torch.empty_like(prob)
    check(_lib().b200_softmax_bwd_rows(p(prob), p(dprob), p(out), c_int(_DT[prob.dtype]), c_longlong(rows), c_int(cols),
                                       c_longlong(cols), c_float(scale), stream_ptr()), "softmax_bwd_rows")
    _count()
    return out


# ------------------------------------------------------------------------------------------------
# multi-tensor AdamW / clipping
# ------------------------------------------------------------------------------------------------
class _AdamWTensor(ctypes.Structure):
    _fields_ = [("p", c_void_p), ("g", c_void_p), ("m", c_void_p), ("v", c_void_p), ("shadow", c_void_p),
                ("n", c_longlong), ("aligned16", c_int), ("pad_", c_int)]


class TensorTable:
    """Device-resident descriptor table over a list of (param, grad, m, v, shadow) tensors."""

    def __init__(self, params, grads, ms=None, vs=None, shadows=None):
        n = len(params)
        chunk = _lib().b200_adamw_chunk_elems()
        arr = (_AdamWTensor * n)()
        starts = []
        total = 0
        for i in range(n):
            ptrs = [params[i].data_ptr() if params[i] is not None else 0,
                    grads[i].data_ptr() if grads[i] is not None else 0,
                    ms[i].data_ptr() if ms is not None else 0,
                    vs[i].data_ptr() if vs is not None else 0]
            sh = shadows[i] if shadows is not None else None
            numel = (grads[i] if grads[i] is not None else params[i]).numel()
            arr[i].p, arr[i].g, arr[i].m, arr[i].v = ptrs
            arr[i].shadow = sh.data_ptr() if sh is not None else 0
            arr[i].n = numel
            al = all(q % 16 == 0 for q in ptrs) and (sh is None or sh.data_ptr() % 8 == 0)
            al32 = all(q % 32 == 0 for q in ptrs) and (sh is None or sh.data_ptr() % 16 == 0)
            arr[i].aligned16 = 2 if al32 else int(al)
            starts.append(total)
            total += (numel + chunk - 1) // chunk
        dev = (grads[0] if grads[0] is not None else params[0]).device
        raw = bytes(arr)
        self.table = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(dev)
        self.starts = torch.tensor(starts, dtype=torch.int64).to(dev)
        self.n = n
        self.total_chunks = total
        self.device = dev


class AdamWLaunch:
    """Holds the table, device hyper-parameters and the device step counter of one fused group."""

    def __init__(self, params, grads, ms, vs, shadows, step0: int, mode: int = 0):
        self.tt = TensorTable(params, grads, ms, vs, shadows)
        # the table stores raw addresses: keep every tensor alive for as long as the launch object exists
        self._keepalive = (list(params), list(grads), list(ms) if ms is not None else None,
                           list(vs) if vs is not None else None, list(shadows) if shadows is not None else None)
        dev = self.tt.device
        self.hyper = torch.zeros(8, dtype=torch.float32, device=dev)
        # ring of pinned staging rows: the host may run several steps ahead of the async copies
        self._hyper_host = torch.zeros(16, 8, dtype=torch.float32).pin_memory()
        self._hyper_slot = 0
        self._hyper_vals = None
        self.step = torch.tensor([step0], dtype=torch.int32, device=dev)
        self.mode = mode

    def set_hyper(self, lr, b1, b2, eps, wd, grad_scale=1.0) -> None:
        vals = (float(lr), float(b1), float(b2), float(eps), float(wd), float(grad_scale))
        if vals == self._hyper_vals:
            return
        if self.hyper.is_cuda and torch.cuda.is_current_stream_capturing():
            # a copy recorded into a CUDA graph would overwrite the refreshed values on every replay with whatever
            # this staging row holds by then; leave the device copy alone (the trainer pushes pending values with
            # refresh_device_hyper() right before capturing and before every replay)
            return
        self._hyper_vals = vals
        row = self._hyper_host[self._hyper_slot]
        self._hyper_slot = (self._hyper_slot + 1) % self._hyper_host.shape[0]
        row[:6] = torch.tensor(vals, dtype=torch.float32)
        self.hyper.copy_(row, non_blocking=True)

    def launch(self, max_blocks: int = 0) -> None:
        """``max_blocks`` caps the grid (0 = fill the GPU); a small cap keeps SM room for concurrent GEMMs."""
        check(_lib().b200_adamw_multi_grid(p(self.tt.table), p(self.tt.starts), c_int(self.tt.n),
                                           c_longlong(self.tt.total_chunks), p(self.hyper), p(self.step),
                                           c_int(self.mode), c_int(1), c_int(max_blocks), stream_ptr()),
              "adamw_multi")
        _count(2)
        bump_param_epoch()


class DeviceHyper:
    """AdamW hyper-parameters (lr, beta1, beta2, eps, weight_decay, grad_scale) and the step counter ``t`` in device
    memory, for kernels that must stay valid inside a CUDA graph: values change through a pinned staging ring and an
    async copy, never through a kernel argument."""

    def __init__(self, device, step0: int = 1):
        self.hyper = torch.zeros(8, dtype=torch.float32, device=device)
        self._host = torch.zeros(16, 8, dtype=torch.float32).pin_memory()
        self._slot = 0
        self._vals = None
        self.step = torch.tensor([step0], dtype=torch.int32, device=device)

    def set(self, lr, b1, b2, eps, wd, grad_scale=1.0) -> None:
        vals = (float(lr), float(b1), float(b2), float(eps), float(wd), float(grad_scale))
        if vals == self._vals or torch.cuda.is_current_stream_capturing():
            return  # (a copy recorded into a graph would replay stale staging contents: see AdamWLaunch.set_hyper)
        self._vals = vals
        row = self._host[self._slot]
        self._slot = (self._slot + 1) % self._host.shape[0]
        row[:6] = torch.tensor(vals, dtype=torch.float32)
        self.hyper.copy_(row, non_blocking=True)

    def advance(self) -> None:
        check(_lib().b200_incr_step(p(self.step), stream_ptr()), "incr_step")
        _count()


class _WgradAdamwArgs(ctypes.Structure):
    _fields_ = [("p", c_void_p * 3), ("m", c_void_p * 3), ("v", c_void_p * 3), ("rows", c_int * 3), ("nseg", c_int),
                ("mode", c_int), ("shadow", c_void_p), ("hyper", c_void_p), ("step", c_void_p)]


def wgrad_adamw(dy: Tensor, x: Tensor, params, ms, vs, shadow: Tensor | None, dh: DeviceHyper, mode: int = 0,
                grad_out: Tensor | None = None, pair: bool | None = None) -> None:
    """``dW = dyᵀ·x`` consumed by AdamW inside the GEMM epilogue (``EPI_ADAMW``): ``params`` (1..3 contiguous fp32
    ``[rows_i, N]`` tensors stacked along the rows of dW, in the order of their fused forward GEMM) and their moments
    are updated in place, ``shadow`` (bf16 ``[Σrows, N]``) is rewritten; the gradient is only stored when ``grad_out``
    (fp32 ``[Σrows, N]``) is given."""
    T, M = dy.shape
    N = x.shape[1]
    assert x.shape[0] == T and dy.dtype == x.dtype == torch.bfloat16 and dy.stride(1) == 1 and x.stride(1) == 1
    assert 1 <= len(params) <= 3 and sum(q.shape[0] for q in params) == M
    a = _WgradAdamwArgs()
    for i, (q, m_, v_) in enumerate(zip(params, ms, vs)):
        assert q.dtype == torch.float32 and q.is_contiguous() and q.shape[1] == N
        assert m_.is_contiguous() and v_.is_contiguous() and m_.shape == q.shape == v_.shape
        a.p[i], a.m[i], a.v[i], a.rows[i] = q.data_ptr(), m_.data_ptr(), v_.data_ptr(), q.shape[0]
    a.nseg, a.mode = len(params), mode
    if shadow is not None:
        assert shadow.dtype == torch.bfloat16 and shadow.is_contiguous() and tuple(shadow.shape) == (M, N)
        a.shadow = shadow.data_ptr()
    a.hyper, a.step = dh.hyper.data_ptr(), dh.step.data_ptr()
    if grad_out is not None:
        assert grad_out.dtype == torch.float32 and grad_out.is_contiguous() and tuple(grad_out.shape) == (M, N)
    # the kernel computes the transpose (rows = N input features, columns = the M stacked parameter rows)
    if pair is None:
        use_pair, bn = _plan_gemm(N, M, True, N >= 256) if GEMM_PAIR == "auto" else (GEMM_PAIR == "1" and N >= 256, 0)
    else:
        use_pair, bn = pair and N >= 256, 0
    check(_lib().b200_wgrad_adamw(p(dy), p(x), c_int(T), c_int(M), c_int(N), c_int(dy.stride(0)), c_int(x.stride(0)),
                                  ctypes.byref(a), p(grad_out), c_int(int(use_pair)), c_int(bn), stream_ptr()),
          "wgrad_adamw")
    _count()


_clip_plans: dict = {}


def clip_grad_norm(grads: list[Tensor], max_norm: float, eps: float, table: TensorTable | None = None) -> Tensor:
    """Global-L2 clip of contiguous fp32 ``grads`` in place (3 launches).  The descriptor table, the partial-sum
    buffer and the result are built once per set of gradient addresses and reused (arena-backed gradients keep
    their addresses), so steady-state calls neither allocate nor copy from the host — which is also what lets a
    clipped step be captured in a CUDA graph."""
    key = tuple((g.data_ptr(), g.numel()) for g in grads)
    plan = None if table is not None else _clip_plans.get(key)
    if plan is None:
        if table is None and torch.cuda.is_current_stream_capturing():
            raise RuntimeError("clip_grad_norm: first call for this set of gradients must happen before stream "
                               "capture (it builds the descriptor table with a host-to-device copy)")
        tt = table or TensorTable([None] * len(grads), grads)
        nparts = _lib().b200_gradnorm_partials(c_longlong(tt.total_chunks))
        plan = (tt, torch.empty(nparts, dtype=torch.float32, device=tt.device),
                torch.empty(2, dtype=torch.float32, device=tt.device), list(grads))
        if table is None:
            if len(_clip_plans) >= 8:
                _clip_plans.pop(next(iter(_clip_plans)))
            _clip_plans[key] = plan
    tt, partial, out, _keep = plan
    check(_lib().b200_clip_grad_norm(p(tt.table), p(tt.starts), c_int(tt.n), c_longlong(tt.total_chunks),
                                     c_float(max_norm), c_float(eps), p(partial), p(out), c_int(1), stream_ptr()),
          "clip_grad_norm")
    _count(3)
    return out[0]


# ------------------------------------------------------------------------------------------------
# FlashAttention-2
# ------------------------------------------------------------------------------------------------
def _strides3(t: Tensor):
    assert t.dim() == 4 and t.stride(3) == 1, "attention operands are (B, H, N, D) views with unit inner stride"
    return (c_longlong * 3)(t.stride(0), t.stride(1), t.stride(2))


def cast3_bf16(a: Tensor, b: Tensor | None = None, c: Tensor | None = None):
    """bf16 copies of up to three contiguous fp32 tensors in ONE launch."""
    srcs = [t for t in (a, b, c) if t is not None]
    for t in srcs:
        assert t.dtype == torch.float32 and t.is_contiguous()
    outs = [torch.empty(t.shape, dtype=torch.bfloat16, device=t.device) for t in srcs]
    args = []
    for i in range(3):
        if i < len(srcs):
            args += [p(srcs[i]), p(outs[i]), c_longlong(srcs[i].numel())]
        else:
            args += [c_void_p(0), c_void_p(0), c_longlong(0)]
    check(_lib().b200_cast3_f32_to_bf16(*args, stream_ptr()), "cast3")
    _count()
    return outs


def fa2_fwd(q: Tensor, k: Tensor, v: Tensor, causal: bool, out: Tensor | None = None, out_f32: bool = False,
            cluster: bool = True):
    """``q,k,v``: bf16 ``(B, H, N, D)`` views (any batch/head/row strides).  Returns ``(o, lse)``;
    ``o`` has q's layout unless ``out`` is given (bf16, or fp32 with ``out_f32``), ``lse`` is fp32 ``(B, H, Nq)``.
    ``cluster``: let pairs of CTAs share K/V tiles through TMA multicast when the query-tile count is even."""
    B, H, Nq, D = q.shape
    Nk = k.shape[2]
    assert q.dtype == k.dtype == v.dtype == torch.bfloat16
    odt = torch.float32 if out_f32 else torch.bfloat16
    if out is None:
        out = torch.empty_strided(q.shape, q.stride(), dtype=odt, device=q.device) \
            if q.is_contiguous() else torch.empty(B, Nq, H, D, dtype=odt, device=q.device).permute(0, 2, 1, 3)
    assert out.dtype == odt
    lse = torch.empty(B, H, Nq, dtype=torch.float32, device=q.device)
    check(_lib().b200_fa2_fwd_ex(p(q), p(k), p(v), p(out), p(lse), c_int(B), c_int(H), c_int(Nq), c_int(Nk), c_int(D),
                                 _strides3(q), _strides3(k), _strides3(v), _strides3(out), c_int(int(causal)),
                                 c_float(1.0 / (D ** 0.5)), c_int(int(out_f32)), c_int(int(cluster)), stream_ptr()),
          "fa2_fwd")
    _count()
    return out, lse


_dq_workspaces: dict = {}


def fa2_bwd(q, k, v, o, lse, do, causal: bool, dq=None, dk=None, dv=None, rope_cache: Tensor | None = None,
            grads_f32: bool = False):
    """``rope_cache``: when q,k were rotated before attention, un-rotate dq,dk on the way out (fused).
    ``o`` / ``do`` may be fp32 (the delta pre-pass reads them as they are and emits the bf16 ``do`` the tensor cores
    need); ``grads_f32`` makes the kernels write fp32 dq/dk/dv."""
    B, H, Nq, D = q.shape
    Nk = k.shape[2]
    if do.stride(3) != 1:
        do = do.contiguous()
    gdt = torch.float32 if grads_f32 else q.dtype
    dq = torch.empty(q.shape, dtype=gdt, device=q.device) if dq is None else dq
    dk = torch.empty(k.shape, dtype=gdt, device=q.device) if dk is None else dk
    dv = torch.empty(v.shape, dtype=gdt, device=q.device) if dv is None else dv
    delta = torch.empty(B, H, Nq, dtype=torch.float32, device=q.device)
    # fp32 dQ accumulator: one zeroed workspace per shape, handed back zeroed by the kernel's last pass
    # (calls of the same shape are assumed to be stream-ordered, as they are in a training step)
    key = (q.device, B, H, Nq, D)
    dq_acc = _dq_workspaces.get(key)
    if dq_acc is None:
        dq_acc = torch.zeros(B, H, Nq, D, dtype=torch.float32, device=q.device)
        _dq_workspaces[key] = dq_acc
    rope_ctx = 0
    if rope_cache is not None:
        assert rope_cache.dtype == torch.float32 and rope_cache.is_contiguous() and rope_cache.shape[2] * 2 == D
        rope_ctx = rope_cache.shape[1]
    flags = 1 | (2 if o.dtype == torch.float32 else 0) | (4 if do.dtype == torch.float32 else 0) | (8 if grads_f32 else 0)
    do_scratch = torch.empty(B, H, Nq, D, dtype=torch.bfloat16, device=q.device) if do.dtype == torch.float32 else None
    check(_lib().b200_fa2_bwd_ex(p(q), p(k), p(v), p(o), p(lse), p(do), p(dq), p(dk), p(dv), p(delta), p(dq_acc),
                                   c_int(B), c_int(H), c_int(Nq), c_int(Nk), c_int(D), _strides3(q), _strides3(k),
                                   _strides3(v), _strides3(o), _strides3(do), _strides3(dq), _strides3(dk),
                                   _strides3(dv), c_int(int(causal)), c_float(1.0 / (D ** 0.5)), p(rope_cache),
                                   c_int(rope_ctx), c_int(flags), p(do_scratch), stream_ptr()), "fa2_bwd")
    _count(4)
    return dq, dk, dv


# ------------------------------------------------------------------------------------------------
# NVLink collectives
# ------------------------------------------------------------------------------------------------
import os as _os

_MULTIMEM_AUTO = _os.environ.get("B200SYS_MULTIMEM", "1") == "1"

def host_ptr_array(ptrs):
    """ctypes array of device addresses (the C ABI takes peer tables as small host arrays)."""
    return (ctypes.c_ulonglong * len(ptrs))(*[int(x) for x in ptrs])


def allreduce_avg(arena, byte_off: int, numel: int, dtype: torch.dtype, one_shot_max: int, blocks: int,
                  algo: int = 0, signal_slot: int = 0) -> None:
    """In-place average over ranks of ``numel`` elements at ``byte_off`` of the symmetric arena."""
    es = 2 if dtype == torch.bfloat16 else 4
    if algo == 0:
        # measured on B200/NVSwitch (profiles/allreduce_*.json): one-shot only wins for tiny messages; with two
        # ranks the P2P two-shot beats the NVLS path, with more ranks the in-switch reduction moves fewer bytes
        if numel * es <= one_shot_max:
            algo = 1
        elif arena.multicast_ptr and arena.world_size > 2 and _MULTIMEM_AUTO:
            algo = 3
        else:
            algo = 2
    check(_lib().b200_allreduce(arena.peer_ptrs_c, arena.signal_ptrs_c, c_void_p(arena.multicast_ptr),
                                c_longlong(byte_off), c_longlong(numel), c_int(_DT[dtype]),
                                c_float(1.0 / arena.world_size), c_int(arena.rank), c_int(arena.world_size),
                                c_int(algo), c_int(blocks), c_int(signal_slot), stream_ptr()), "allreduce")
    _count()


def bcast_pull(arena, byte_off: int, nbytes: int, src: int, signal_slot: int = 0) -> None:
    """Every rank but ``src`` copies ``nbytes`` at ``byte_off`` of ``src``'s arena into its own (one kernel)."""
    check(_lib().b200_bcast_pull(arena.peer_ptrs_c, arena.signal_ptrs_c, c_longlong(byte_off), c_longlong(nbytes),
                                 c_int(src), c_int(arena.rank), c_int(arena.world_size), c_int(0),
                                 c_int(signal_slot), stream_ptr()), "bcast_pull")
    _count()


def bcast_mc(arena, byte_off: int, nbytes: int, src: int, signal_slot: int = 0) -> None:
    """Rank ``src`` stores ``nbytes`` at ``byte_off`` of its arena into every rank's arena through the NVLS multicast
    mapping (one kernel; needs ``arena.multicast_ptr``)."""
    check(_lib().b200_bcast_mc(arena.peer_ptrs_c, arena.signal_ptrs_c, c_void_p(arena.multicast_ptr),
                               c_longlong(byte_off), c_longlong(nbytes), c_int(src), c_int(arena.rank),
                               c_int(arena.world_size), c_int(0), c_int(signal_slot), stream_ptr()), "bcast_mc")
    _count()


def comm_status() -> tuple | None:
    """None when the NVLink kernels on the current device are healthy; else ``(rank, cta, peer, phase, slot, flag)`` of
    the first cross-rank barrier that ran out of time (phase 1: raising the peer's flag, 2: waiting for the peer)."""
    info = (c_int * 6)()
    rc = _lib().b200_comm_status(info)
    if rc == 0:
        return None
    if rc != 1:
        check(rc, "comm_status")
    return tuple(int(x) for x in info)


def comm_check() -> None:
    st = comm_status()
    if st is not None:
        raise RuntimeError("NVLink communicator dead: rank %d CTA %d gave up on peer %d (phase %d, signal slot %d, "
                           "flag %d) — a rank died, fell behind by more than the barrier time limit, or ranks "
                           "disagree on the kernel sequence" % st)


def comm_reset(timeout_ms: int = 0) -> None:
    """Clear the health block and (``timeout_ms`` > 0) set the cross-rank barrier time limit on the current device."""
    check(_lib().b200_comm_reset(c_longlong(int(timeout_ms))), "comm_reset")


def symm_barrier(arena, signal_slot: int) -> None:
    check(_lib().b200_symm_barrier(arena.signal_ptrs_c, c_int(arena.rank), c_int(arena.world_size),
                                   c_int(signal_slot), stream_ptr()), "symm_barrier")
    _count()


def adamw_multi_distribute(launch: "AdamWLaunch", param_arena, shadow_arena) -> None:
    """Owner AdamW over ``launch``'s table + P2P store of the new values into every rank's arenas."""
    check(_lib().b200_adamw_multi_distribute(
        p(launch.tt.table), p(launch.tt.starts), c_int(launch.tt.n), c_longlong(launch.tt.total_chunks),
        p(launch.hyper), p(launch.step), c_int(launch.mode), c_int(1), param_arena.peer_ptrs_c,
        shadow_arena.peer_ptrs_c if shadow_arena is not None else None,
        ctypes.c_ulonglong(param_arena.bytes.data_ptr()),
        ctypes.c_ulonglong(shadow_arena.bytes.data_ptr() if shadow_arena is not None else 0),
        c_int(param_arena.world_size), stream_ptr()), "adamw_multi_distribute")
    _count(2)
    bump_param_epoch()
