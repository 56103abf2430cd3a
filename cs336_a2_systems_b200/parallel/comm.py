"""Gradient-reduction backends behind the DDP wrappers.

``C10dReducer``  — ``torch.distributed`` all-reduce (NCCL with ``ReduceOp.AVG`` so the average is
    fused, or gloo SUM followed by one in-place scale of the flat buffer).  This is what the
    reference does for every collective and is our fallback + measured baseline.
``SymmReducer``  — (``parallel/symm.py``) buckets in NVLink symmetric memory reduced by our own
    sm_100a kernels with the 1/world_size average fused, launched on a side stream.

Selection: explicit ``backend`` argument, else ``B200SYS_COMM`` in {auto, symm, c10d}; ``auto`` picks
``symm`` only when it is legal (NCCL process group, one distinct P2P-capable GPU per rank).
"""

from __future__ import annotations

import os

import torch
import torch.distributed as dist


class C10dReducer:
    kind = "c10d"

    def __init__(self, items, group, per_parameter: bool, dtypes=None, devices=None):
        self.group = group
        self.world_size = dist.get_world_size(group)
        self.per_parameter = per_parameter
        self.slots = None
        be = dist.get_backend(group)
        self._avg = be == "nccl"  # NCCL fuses the average; gloo has no AVG
        self._bufs: list[torch.Tensor] = []
        if not per_parameter:
            for n, dt, dev in zip(items, dtypes, devices):
                self._bufs.append(torch.zeros(n, dtype=dt, device=dev))

    def buffer(self, bi: int) -> torch.Tensor:
        return self._bufs[bi]

    def _reduce(self, t: torch.Tensor):
        op = dist.ReduceOp.AVG if (self._avg and t.is_cuda) else dist.ReduceOp.SUM
        work = dist.all_reduce(t, op=op, group=self.group, async_op=True)
        return [work, t, op != dist.ReduceOp.SUM]  # mutable: stream_after() may apply the average early

    def reduce_bucket_async(self, bi: int):
        return self._reduce(self._bufs[bi])

    def reduce_param_async(self, idx: int, param: torch.Tensor):
        return self._reduce(param.grad)

    def stream_after(self, handle):
        """A side stream that has been made to wait for this reduction (and applies the gloo-style average)."""
        work, t, averaged = handle
        if not t.is_cuda:
            raise RuntimeError("bucket callbacks need CUDA tensors")
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=t.device)
        with torch.cuda.stream(self._side):
            work.wait()
            if not averaged:
                t.div_(self.world_size)
                handle[2] = True
        return self._side

    def wait(self, handles) -> None:
        for work, _t, _ in handles:
            work.wait()
        for _work, t, averaged in handles:
            if not averaged:
                t.div_(self.world_size)


def _choice(backend: str | None) -> str:
    return (backend or os.environ.get("B200SYS_COMM", "auto")).lower()


def make_reducer(items, group, backend: str | None, per_parameter: bool, dtypes=None, devices=None):
    """``items``: bucket sizes in elements (bucketed) or the parameter list (per-parameter)."""
    choice = _choice(backend)
    if per_parameter:
        dtypes = [p.dtype for p in items]
        devices = [p.device for p in items]
    if choice in ("auto", "symm"):
        from . import symm

        ok, why = symm.is_legal(devices, dtypes, group)
        if ok:
            return symm.SymmReducer(items, group, per_parameter, dtypes, devices)
        if choice == "symm":
            raise RuntimeError(f"B200SYS_COMM=symm requested but the NVLink path is not legal: {why}")
    return C10dReducer(items, group, per_parameter, dtypes, devices)
