"""Device-side timing helpers: CUDA events on the launching stream, warm-up, L2 flush between
iterations, max over ranks.  (The reference uses ``timeit`` + ``cuda.synchronize`` and
``triton.testing.do_bench``)"""

from __future__ import annotations

import statistics
import time
from typing import Callable

import torch
import torch.distributed as dist

_flush_buf: dict = {}


def flush_l2(device=None) -> None:
    """Overwrite a 256 MiB buffer (> the 126 MB L2) so the next kernel starts cold."""
    dev = torch.device(device or torch.cuda.current_device())
    buf = _flush_buf.get(dev)
    if buf is None:
        buf = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device=dev)
        _flush_buf[dev] = buf
    buf.zero_()


def time_cuda(fn: Callable[[], object], warmup: int = 3, iters: int = 10, flush: bool = True) -> dict:
    """Per-iteration device times of ``fn`` in ms (CUDA events; synchronised on both sides)."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    times = []
    for _ in range(iters):
        if flush:
            flush_l2()
        s = torch.cuda.Event(enable_timing=True)
        e = torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        times.append(s.elapsed_time(e))
    return summarize(times)


def time_host(fn: Callable[[], object], warmup: int = 1, iters: int = 5) -> dict:
    for _ in range(warmup):
        fn()
    times = []
    for _ in range(iters):
        t0 = time.perf_counter()
        fn()
        times.append((time.perf_counter() - t0) * 1e3)
    return summarize(times)


def summarize(times_ms) -> dict:
    times_ms = list(times_ms)
    return {
        "mean_ms": statistics.fmean(times_ms),
        "median_ms": statistics.median(times_ms),
        "min_ms": min(times_ms),
        "max_ms": max(times_ms),
        "std_ms": statistics.pstdev(times_ms) if len(times_ms) > 1 else 0.0,
        "n": len(times_ms),
    }


def max_over_ranks(value: float, device=None) -> float:
    """A multi-GPU time is the slowest rank's time."""
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    dev = device or (torch.device("cuda", torch.cuda.current_device()) if dist.get_backend() == "nccl" else "cpu")
    t = torch.tensor([value], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def time_collective(fn: Callable[[], object], warmup: int, iters: int, dev, min_seconds: float = 0.25) -> float:
    """Per-call device time (ms) of a multi-rank ``fn``: CUDA events around batches of ``iters`` calls, batches
    repeated until ``min_seconds`` of device time have been timed (long enough for ``nvidia-smi`` to sample the
    clocks of the region), median batch, max over ranks."""
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize(dev)
    if dist.is_available() and dist.is_initialized():
        dist.barrier()
    per_call, total = [], 0.0
    while total < min_seconds * 1e3 and len(per_call) < 50:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        torch.cuda.synchronize(dev)
        ms = e0.elapsed_time(e1)
        per_call.append(ms / iters)
        # every rank must run the same number of batches of the same size: agree on the elapsed time
        agreed = max_over_ranks(ms, dev)
        total += agreed
        if agreed < 5.0:
            iters *= 4  # microsecond-scale calls: grow the batch instead of the batch count
    return max_over_ranks(statistics.median(per_call), dev)
