"""NVTX ranges that degrade to no-ops without CUDA (reference: ``torch.cuda.nvtx`` ranges around
backward / communication / model sub-blocks)."""

from __future__ import annotations

import contextlib
import functools

import torch

_enabled = torch.cuda.is_available()


@contextlib.contextmanager
def range(name: str):  # noqa: A001 - mirrors torch.cuda.nvtx.range
    if _enabled:
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


def annotate(name: str):
    """Decorator form: ``@nvtx.annotate("attention")``."""
    def deco(fn):
        @functools.wraps(fn)
        def wrapped(*a, **k):
            with range(name):
                return fn(*a, **k)
        return wrapped
    return deco
