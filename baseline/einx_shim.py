"""Baseline-only stand-in for the three ``einx`` calls of the course model.

``einx`` is not installable offline.  The reference's shipped ``cs336_basics/model.py`` uses it at
lines 141, 146, 501, 511-512 with five fixed patterns; this module implements exactly those so the
UNMODIFIED reference model can run in the ``--impl reference`` arm of ``bench.py``.  Never imported
by the product package.
"""
import torch


def get_at(pattern, tensor, index):
    if pattern.replace(" ", "") != "cos_sin[pos]half_dim,...->cos_sin...half_dim":
        raise NotImplementedError(pattern)
    return tensor[:, index, :]


def rearrange(pattern, *tensors, **kw):
    key = " ".join(pattern.split())
    if key == "... x_half, ... x_half -> ... (x_half (1 + 1))":
        a, b = tensors
        return torch.stack((a, b), dim=-1).flatten(-2)
    (t,) = tensors
    lead = [1] * len(kw.get("b", []))
    if key == "seq -> b... seq":
        return t.view(*lead, t.shape[0])
    if key == "query -> b... 1 query 1":
        return t.view(*lead, 1, t.shape[0], 1)
    if key == "key -> b... 1 1 key":
        return t.view(*lead, 1, 1, t.shape[0])
    raise NotImplementedError(pattern)
