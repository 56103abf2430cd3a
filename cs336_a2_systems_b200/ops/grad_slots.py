"""Gradient destination registry ("arena-backed gradients").

The reference flattens a bucket's gradients with ``torch.cat`` before each all-reduce and
re-materialises ``param.grad`` from the reduced buffer afterwards
(``ddp_bucketed_overlapped_sharded.py:297,311-314``).  Here whoever owns a flat gradient
buffer (a DDP bucket in NVLink symmetric memory, a flat optimizer arena) registers, per
parameter, the view its gradient should be *produced into*.  The weight-gradient kernels
(tcgen05 wgrad GEMM epilogue, RMSNorm/embedding backward) look the slot up and write there
directly, and autograd's AccumulateGrad adopts the returned view as ``param.grad`` without a
copy — flatten/unflatten vanish.

A slot can be used once per backward (``claim``); a second producer for the same parameter
(tied weights, gradient accumulation) gets ``None`` and falls back to an ordinary temporary
that autograd accumulates, so semantics never change.
"""

from __future__ import annotations

import torch
from torch.utils.weak import WeakIdKeyDictionary

# identity-keyed (tensor == is elementwise, so an ordinary WeakKeyDictionary cannot hold tensors)
_slots = WeakIdKeyDictionary()


class _Slot:
    __slots__ = ("base", "offset", "shape", "claimed")

    def __init__(self, base: torch.Tensor, offset: int, shape):
        self.base, self.offset, self.shape, self.claimed = base, offset, tuple(shape), False

    def view(self) -> torch.Tensor:
        n = 1
        for s in self.shape:
            n *= s
        return self.base.narrow(0, self.offset, n).view(self.shape)


def register(param: torch.Tensor, base: torch.Tensor, offset: int) -> None:
    """``param``'s gradient lives at ``base[offset : offset+numel]`` (a 1-D flat buffer)."""
    _slots[param] = _Slot(base, offset, param.shape)


def unregister(param: torch.Tensor) -> None:
    _slots.pop(param, None)


def claim(param: torch.Tensor) -> torch.Tensor | None:
    """Fresh view of the parameter's slot for a producer to write into, or None."""
    s = _slots.get(param)
    if s is None or s.claimed or param.grad is not None:
        return None
    s.claimed = True
    return s.view()


def slot_view(param: torch.Tensor) -> torch.Tensor | None:
    s = _slots.get(param)
    return None if s is None else s.view()


def release_all(params=None) -> None:
    """Start of a new step: every slot may be produced into again."""
    if params is None:
        for s in _slots.values():
            s.claimed = False
    else:
        for p in params:
            s = _slots.get(p)
            if s is not None:
                s.claimed = False


def claim_group(params) -> torch.Tensor | None:
    """One ``[Σ rows, cols]`` view covering the slots of several 2-D parameters, if they are all unclaimed
    and lie back to back in the same flat buffer in this order (then ONE wgrad GEMM can produce all of
    them); otherwise None and nothing is claimed."""
    slots = [_slots.get(p) for p in params]
    if any(s is None or s.claimed for s in slots) or any(p.grad is not None for p in params):
        return None
    base, off0 = slots[0].base, slots[0].offset
    cols = params[0].shape[1]
    off = off0
    rows = 0
    for p, s in zip(params, slots):
        if s.base is not base or s.offset != off or p.dim() != 2 or p.shape[1] != cols:
            return None
        off += p.numel()
        rows += p.shape[0]
    for s in slots:
        s.claimed = True
    return base.narrow(0, off0, rows * cols).view(rows, cols)


def fusion_friendly_order(model) -> list:
    """Parameters of ``model`` in an order where the members of every fused GEMM group are adjacent: q/k/v already
    are (registration order); for SwiGLU blocks (``w1, w2, w3``) the up-projection ``w3`` is moved next to the gate
    ``w1`` so that :func:`claim_group` can hand the fused w1‖w3 wgrad GEMM one joint slot."""
    params = list(model.parameters())
    pos = {id(p): i for i, p in enumerate(params)}
    for m in model.modules():
        w1, w2, w3 = (getattr(getattr(m, n, None), "weight", None) for n in ("w1", "w2", "w3"))
        if w1 is None or w2 is None or w3 is None:
            continue
        i1, i2, i3 = pos.get(id(w1)), pos.get(id(w2)), pos.get(id(w3))
        if i1 is None or i2 is None or i3 is None or not (i2 == i1 + 1 and i3 == i2 + 1):
            continue
        params[i2], params[i3] = w3, w2
        pos[id(w3)], pos[id(w2)] = i2, i3
    return params
