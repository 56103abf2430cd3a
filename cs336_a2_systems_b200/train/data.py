"""Batching utilities: random-window language-model batches and synthetic token streams.

``get_batch`` mirrors ``cs336_basics/data.py:10-30`` (random windows of a token array, targets shifted
by one, pinned + non-blocking copies for CUDA).  ``SyntheticBatches`` is what the drivers and benchmarks
use when no dataset is given (the reference generated ``data_large.pkl`` / ``random_data.pt`` offline).
"""
from __future__ import annotations

import numpy as np
import torch


def get_batch(dataset, batch_size: int, context_length: int, device: str | torch.device):
    starts = torch.randint(len(dataset) - context_length, (batch_size,))
    x = torch.stack([torch.from_numpy(np.asarray(dataset[i:i + context_length]).astype(np.int64)) for i in starts])
    y = torch.stack([torch.from_numpy(np.asarray(dataset[i + 1:i + 1 + context_length]).astype(np.int64)) for i in starts])
    if "cuda" in str(device):
        return x.pin_memory().to(device, non_blocking=True), y.pin_memory().to(device, non_blocking=True)
    return x.to(device), y.to(device)


def set_seed(seed: int) -> None:
    """Seed python-side generators the way the reference drivers do (``naive_ddp.py:14-20``)."""
    import random
    random.seed(seed)
    np.random.seed(seed)
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)


def generate_random_data(num_batches: int, batch_size: int, context_length: int, vocab_size: int, seed: int = 0,
                         path: str | None = None):
    """``(inputs, targets)`` int64 tensors of shape ``(num_batches*batch_size, ctx)``; optionally saved with
    ``torch.save`` (reference ``naive_ddp.py:22-26`` wrote ``random_data.pt``)."""
    g = torch.Generator().manual_seed(seed)
    x = torch.randint(0, vocab_size, (num_batches * batch_size, context_length), generator=g)
    y = torch.roll(x, -1, dims=1)
    if path:
        torch.save((x, y), path)
    return x, y


class SyntheticBatches:
    """Endless pinned-host synthetic batches; ``rank``/``world_size`` shard the global batch like the
    reference's ``batch_full[rank*bs_p:(rank+1)*bs_p]`` (``ddp_bucketed_overlapped_sharded.py:163-166``)."""

    def __init__(self, global_batch: int, context_length: int, vocab_size: int, rank: int = 0, world_size: int = 1,
                 seed: int = 0, pool: int = 16, pin: bool | None = None):
        assert global_batch % world_size == 0
        self.local = global_batch // world_size
        g = torch.Generator().manual_seed(seed)
        pin = torch.cuda.is_available() if pin is None else pin
        self.xs, self.ys = [], []
        for _ in range(pool):
            full = torch.randint(0, vocab_size, (global_batch, context_length), generator=g)
            x = full[rank * self.local:(rank + 1) * self.local].contiguous()
            y = torch.roll(x, -1, dims=1)
            self.xs.append(x.pin_memory() if pin else x)
            self.ys.append(y.pin_memory() if pin else y)
        self._i = 0

    def __iter__(self):
        return self

    def __next__(self):
        i = self._i % len(self.xs)
        self._i += 1
        return self.xs[i], self.ys[i]
