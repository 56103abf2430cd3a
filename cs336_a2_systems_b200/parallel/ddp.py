"""Data-parallel wrappers with communication overlapped with backward.

``DDPIndividualParameters`` — one asynchronous gradient all-reduce per parameter, issued from a
    post-accumulate-grad hook (reference ``DDP``, ``ddp_bucketed_overlapped_sharded.py:217-248``).
``DDPBucketed`` — reverse-order greedy buckets of at most ``bucket_size_mb``; a bucket is reduced
    as soon as its last gradient lands (reference ``DDP_Bucketed``, ``:251-318``).

Both hold the wrapped model as the registered submodule ``module`` and add no parameters, so
``ddp.named_parameters()`` lines up with the bare model's and ``ddp.parameters()`` feeds the
optimizer (reference ``tests/test_ddp.py:86-89``).

B200-first design
-----------------
* Gradients LIVE in the bucket: every trainable parameter's ``.grad`` is a view into its
  bucket's flat buffer (``ops.grad_slots``), so there is no flatten before and no unflatten
  after the reduction, and the 1/world_size average is fused into the reduction itself.
* Two communication backends picked once at construction (``parallel.comm``):
  ``symm`` — buckets are carved out of an NVLink symmetric-memory arena and reduced by our own
  sm_100a kernels (one-shot / two-shot P2P or NVLS multimem by message size) on a side stream
  ordered by events, never a host sync; ``c10d`` — ``torch.distributed`` (NCCL or gloo),
  used whenever the NVLink path is not legal (CPU tensors, gloo, several ranks on one GPU)
  and as the measured baseline.
* Initial parameter sync is ONE broadcast of a packed buffer per dtype instead of one per tensor.

Departures from the reference's quirks: frozen parameters are broadcast but never
bucketed or reduced; a trainable parameter that received no gradient does not crash ``finish``;
tied weights are hooked once.
"""

from __future__ import annotations

import os

from collections import OrderedDict

import torch
import torch.distributed as dist
import torch.nn as nn

from ..ops import grad_slots
from ..utils import nvtx
from . import comm as _comm
from .process_group import world


def _unique_params(module: nn.Module):
    seen = set()
    for p in module.parameters():
        if id(p) not in seen:
            seen.add(id(p))
            yield p


@torch.no_grad()
def broadcast_parameters(module: nn.Module, src: int = 0, group=None, buffers: bool = False,
                         backend: str | None = None) -> None:
    """Make every rank's parameters equal to ``src``'s with one packed broadcast per (dtype, device).
    ``backend``: same choice as the gradient reducer's (None = ``B200SYS_COMM``)."""
    _, ws = world()
    if ws == 1:
        return
    tensors = list(_unique_params(module))
    if buffers:
        tensors += [b for b in module.buffers()]
    groups: "OrderedDict[tuple, list[torch.Tensor]]" = OrderedDict()
    for t in tensors:
        groups.setdefault((t.dtype, t.device), []).append(t.data)
    if _comm._choice(backend) in ("auto", "symm"):
        # NVLink path: one pull kernel over a packed symmetric buffer
        from . import symm
        ok, _ = symm.is_legal([t.device for t in tensors], [torch.float32] * len(tensors), group)
        if ok and all(t.is_contiguous() for t in tensors):
            symm.broadcast_tensors([t.data for t in tensors], src, group)
            return
    for (_dtype, _dev), ts in groups.items():
        flat = torch.cat([t.reshape(-1) for t in ts])
        dist.broadcast(flat, src=src, group=group)
        off = 0
        for t in ts:
            n = t.numel()
            t.copy_(flat[off:off + n].view_as(t))
            off += n


class _DDPBase(nn.Module):
    def __init__(self, module: nn.Module, process_group=None, broadcast_buffers: bool = False,
                 backend: str | None = None):
        super().__init__()
        self.module = module
        self.process_group = process_group
        self.rank, self.world_size = world()
        #: True once an optimizer path that consumes the gradients inside its own reduction has been attached
        #: (ShardedOptimizer.enable_symmetric): ``param.grad`` then holds this rank's local gradient, not the average
        self.grads_are_local = False
        broadcast_parameters(module, 0, process_group, buffers=broadcast_buffers, backend=backend)
        self._trainable = [p for p in _unique_params(module) if p.requires_grad]

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def __getattr__(self, name):  # delegate unknown attributes (e.g. .config, .generate) to the model
        try:
            return super().__getattr__(name)
        except AttributeError:
            return getattr(super().__getattr__("module"), name)


class DDPIndividualParameters(_DDPBase):
    """Per-parameter overlapped DDP.

    The hook issues the averaging all-reduce for a gradient the moment autograd has finished
    accumulating it; :meth:`finish_gradient_synchronization` orders the optimizer after all of them.
    """

    def __init__(self, module: nn.Module, process_group=None, backend: str | None = None):
        super().__init__(module, process_group, backend=backend)
        self._pending: list = []
        self._reducer = None
        if self.world_size > 1:
            self._reducer = _comm.make_reducer(self._trainable, self.process_group, backend,
                                               per_parameter=True)
            if self._reducer.kind == "symm":
                # gradients are produced straight into the symmetric arena (one slot per parameter)
                for p, (base, off) in zip(self._trainable, self._reducer.slots):
                    grad_slots.register(p, base, off)
            for i, p in enumerate(self._trainable):
                p.register_post_accumulate_grad_hook(self._make_hook(i))

    def _make_hook(self, idx: int):
        def hook(param: torch.Tensor):
            if param.grad is None:
                return
            with nvtx.range("ddp.allreduce_param"):
                self._pending.append(self._reducer.reduce_param_async(idx, param))
        return hook

    def finish_gradient_synchronization(self) -> None:
        if self._reducer is None:
            return
        self._reducer.wait(self._pending)
        self._pending.clear()
        grad_slots.release_all(self._trainable)

    def on_train_batch_start(self) -> None:
        grad_slots.release_all(self._trainable)


class _Bucket:
    __slots__ = ("params", "offsets", "numel", "flat", "ready", "expected", "dtype", "device", "launched")

    def __init__(self, dtype, device):
        self.params: list[torch.Tensor] = []
        self.offsets: list[int] = []
        self.numel = 0
        self.flat: torch.Tensor | None = None
        self.ready = 0
        self.expected = 0
        self.dtype, self.device = dtype, device
        self.launched = False


def plan_buckets(params: list[torch.Tensor], bucket_size_mb: float | None) -> list[list[int]]:
    """Greedy bucketing over ``params`` (already in reduction order): a bucket is closed when the
    next tensor would push it past the cap, or when dtype/device changes.  A tensor larger than
    the cap gets a bucket of its own.  Returns lists of indices into ``params``."""
    cap = float("inf") if bucket_size_mb is None else bucket_size_mb * 1024 * 1024
    buckets: list[list[int]] = []
    cur: list[int] = []
    cur_bytes = 0
    cur_key = None
    for i, p in enumerate(params):
        nbytes = p.numel() * p.element_size()
        key = (p.dtype, p.device)
        if cur and (cur_bytes + nbytes > cap or key != cur_key):
            buckets.append(cur)
            cur, cur_bytes = [], 0
        cur.append(i)
        cur_bytes += nbytes
        cur_key = key
    if cur:
        buckets.append(cur)
    return buckets


class DDPBucketed(_DDPBase):
    """Bucketed overlapped DDP with arena-backed gradients.

    ``bucket_size_mb=None`` means a single unbounded bucket.  Call
    :meth:`on_train_batch_start` at the top of a step (optional: ``finish`` also resets) and
    :meth:`finish_gradient_synchronization` after ``backward()`` and before ``optimizer.step()``.
    """

    #: alignment of every gradient slot inside a bucket (256-bit vectors of the streaming optimizer kernel)
    ALIGN_BYTES = 32

    def __init__(self, module: nn.Module, bucket_size_mb: float | None = None, process_group=None,
                 backend: str | None = None):
        super().__init__(module, process_group, backend=backend)
        self.bucket_size_mb = bucket_size_mb
        order = list(reversed(self._trainable))  # gradients become ready roughly in reverse order
        self.buckets: list[_Bucket] = []
        self._where: dict[int, tuple[int, int]] = {}
        # in-bucket slot order: the fusion-friendly order that puts w3 next to w1, so the SwiGLU pair gets ONE wgrad GEMM
        # and — under the bucket-fused optimizer — an arena-resident bf16 shadow instead of a per-step re-cast (the
        # default since round 2: every 1/2/8-GPU number of profiles/r2 was measured with it; B200SYS_FUSE_W13_WGRAD=0
        # restores forward parameter order)
        if os.environ.get("B200SYS_FUSE_W13_WGRAD", "1") == "1":
            slot_pos = {id(p): k for k, p in enumerate(grad_slots.fusion_friendly_order(module))}
        else:
            slot_pos = {id(p): k for k, p in enumerate(module.parameters())}
        for idxs in plan_buckets(order, bucket_size_mb):
            b = _Bucket(order[idxs[0]].dtype, order[idxs[0]].device)
            align = max(1, self.ALIGN_BYTES // order[idxs[0]].element_size())
            # membership follows the (reversed) readiness order, but inside a bucket the slots are laid out in
            # forward parameter order so that q/k/v gradients are adjacent (one wgrad GEMM writes all three)
            for i in sorted(idxs, key=lambda i: slot_pos[id(order[i])]):
                p = order[i]
                b.params.append(p)
                b.offsets.append(b.numel)
                b.numel += (p.numel() + align - 1) // align * align
            b.expected = len(b.params)
            self.buckets.append(b)
        self._reducer = None
        self._pending: list = []
        self._bucket_callback = None
        self._reduce_override = None
        if self.world_size > 1:
            self._reducer = _comm.make_reducer([b.numel for b in self.buckets], self.process_group, backend,
                                               per_parameter=False,
                                               dtypes=[b.dtype for b in self.buckets],
                                               devices=[b.device for b in self.buckets])
        for bi, b in enumerate(self.buckets):
            if self._reducer is not None:
                b.flat = self._reducer.buffer(bi)
            else:
                b.flat = torch.zeros(b.numel, dtype=b.dtype, device=b.device)
            for pi, (p, off) in enumerate(zip(b.params, b.offsets)):
                self._where[id(p)] = (bi, pi)
                grad_slots.register(p, b.flat, off)
                if self.world_size > 1:
                    p.register_post_accumulate_grad_hook(self._hook)

    # ---------------------------------------------------------------------------------------
    def _slot(self, b: _Bucket, pi: int) -> torch.Tensor:
        p = b.params[pi]
        return b.flat.narrow(0, b.offsets[pi], p.numel()).view_as(p)

    def _hook(self, param: torch.Tensor) -> None:
        if param.grad is None:
            return
        bi, pi = self._where[id(param)]
        b = self.buckets[bi]
        slot = self._slot(b, pi)
        g = param.grad
        if g.data_ptr() != slot.data_ptr():
            # gradient was produced outside the arena (generic autograd op): move it in once;
            # from now on .grad aliases the bucket
            slot.copy_(g)
            param.grad = slot
        b.ready += 1
        if b.ready == b.expected and not b.launched:
            self._launch(bi)

    def _launch(self, bi: int) -> None:
        b = self.buckets[bi]
        b.launched = True
        with nvtx.range(f"ddp.allreduce_bucket{bi}"):
            handle = self._reduce_override(bi) if self._reduce_override is not None else \
                self._reducer.reduce_bucket_async(bi)
            self._pending.append(handle)
        if self._bucket_callback is not None:
            # chain work behind this bucket's reduction (e.g. the optimizer update of exactly these parameters)
            self._bucket_callback(bi, b.params, self._reducer.stream_after(handle))

    def set_bucket_reduce_override(self, fn) -> None:
        """``fn(bucket_index)`` replaces the bucket all-reduce (used by the bucket-fused sharded optimizer, whose
        kernel reduces, updates and redistributes in one go on the reducer's stream)."""
        self._reduce_override = fn

    def set_bucket_callback(self, fn) -> None:
        """``fn(bucket_index, params, stream)`` is called right after a bucket's reduction has been enqueued;
        ``stream`` is a CUDA stream on which the reduced gradients are valid once its prior work completes."""
        self._bucket_callback = fn

    def finish_gradient_synchronization(self) -> None:
        if self._reducer is not None:
            # buckets some of whose parameters got no gradient this step: treat the missing ones as
            # zero so every rank still reduces the same set of buffers
            for bi, b in enumerate(self.buckets):
                if b.launched or b.ready == 0:
                    continue
                for pi, p in enumerate(b.params):
                    if p.grad is None:
                        self._slot(b, pi).zero_()
                self._launch(bi)
            self._reducer.wait(self._pending)
            self._pending.clear()
        self._reset()

    def on_train_batch_start(self) -> None:
        self._reset()

    def _reset(self) -> None:
        for b in self.buckets:
            b.ready = 0
            b.launched = False
        grad_slots.release_all(self._trainable)


# ---------------------------------------------------------------------------------------------
# Non-overlapped strategies kept for the benchmarking studies (reference D4 / D5)
# ---------------------------------------------------------------------------------------------

@torch.no_grad()
def allreduce_gradients_naive(module: nn.Module, group=None) -> None:
    """After ``backward()``: one blocking averaging all-reduce per parameter
    (reference ``ddp_bucketed_overlapped_sharded.py:87-89``, ``naive_ddp.py:252-255``)."""
    _, ws = world()
    if ws == 1:
        return
    for p in _unique_params(module):
        if p.grad is not None:
            dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=group)
            p.grad.div_(ws)


@torch.no_grad()
def allreduce_gradients_flat(module: nn.Module, group=None) -> None:
    """After ``backward()``: a single all-reduce of all gradients packed into one buffer per dtype
    (reference ``naive_ddp.py:546-553``)."""
    _, ws = world()
    if ws == 1:
        return
    groups: "OrderedDict[tuple, list[torch.Tensor]]" = OrderedDict()
    for p in _unique_params(module):
        if p.grad is not None:
            groups.setdefault((p.grad.dtype, p.grad.device), []).append(p.grad)
    for gs in groups.values():
        flat = torch.cat([g.reshape(-1) for g in gs])
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.div_(ws)
        off = 0
        for g in gs:
            n = g.numel()
            g.copy_(flat[off:off + n].view_as(g))
            off += n
