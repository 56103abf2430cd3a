"""NVLink 5 / NVSwitch symmetric-memory collectives driven by our own sm_100a kernels.

The reference's only transport is ``torch.distributed``.  Here every rank
allocates its gradient buckets (and, for the fused sharded optimizer, its parameters) from a
*symmetric arena*: the same allocation exists on every GPU, all of them are mapped into every
process (CUDA VMM peer mappings + an NVLS multicast object, set up through
``torch.distributed._symmetric_memory`` — framework-layer rendezvous only), and the kernels in
``csrc/comm_allreduce.cu`` / ``csrc/comm_adamw.cu`` read and write peer memory directly:

* ``one-shot``  (small): every rank loads all peers' copies, reduces in registers in rank order,
  barriers, writes its own copy.  Latency-optimal.
* ``two-shot``  (medium/large without NVLS): rank r reduces slice r from all peers and stores the
  result into every peer's slice r.  2·(ws-1)/ws·S bytes over NVLink per GPU.
* ``multimem``  (NVLS): ``multimem.ld_reduce`` of the owned slice through the switch +
  ``multimem.st`` broadcast — ≈S bytes per GPU.

All three fuse the 1/world_size average (and an optional bf16 cast of a shadow copy) and
synchronise ranks with release/acquire flag words in the arena's signal pad — no host sync,
no NCCL call on the hot path.
"""

from __future__ import annotations

import os

import torch
import torch.distributed as dist

_ONE_SHOT_MAX_BYTES = int(os.environ.get("B200SYS_ONESHOT_MAX", 64 * 1024))
_COMM_BLOCKS = int(os.environ.get("B200SYS_COMM_BLOCKS", 0))  # 0: kernel default (16 small / 48 large)
_ALGO = int(os.environ.get("B200SYS_ALLREDUCE_ALGO", 0))  # 0 auto, 1 one-shot, 2 two-shot, 3 multimem


def is_legal(devices, dtypes, group) -> tuple[bool, str]:
    """The NVLink path needs: NCCL group, CUDA tensors, fp32/bf16, one distinct GPU per rank on one node."""
    if not torch.cuda.is_available():
        return False, "no CUDA"
    if any(torch.device(d).type != "cuda" for d in devices):
        return False, "non-CUDA tensors"
    if any(dt not in (torch.float32, torch.bfloat16) for dt in dtypes):
        return False, "unsupported dtype"
    if len({torch.device(d) for d in devices}) != 1:
        return False, "parameters span several devices"
    if dist.get_backend(group) != "nccl":
        return False, "process group is not NCCL"
    ws = dist.get_world_size(group)
    if ws > torch.cuda.device_count():
        return False, "more ranks than GPUs (ranks share a device)"
    if ws > 8:
        return False, "more than one NVSwitch domain"
    try:
        import importlib

        importlib.import_module("torch.distributed._symmetric_memory")
    except Exception as e:  # pragma: no cover
        return False, f"symmetric memory unavailable: {e}"
    from .. import _native

    if not _native.have_lib():
        return False, "libb200sys.so not built"
    return True, ""


class SymmArena:
    """One symmetric allocation per rank, peer-mapped everywhere, plus its signal pad."""

    def __init__(self, nbytes: int, device, group):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        self.rank = dist.get_rank(self.group)
        self.world_size = dist.get_world_size(self.group)
        nbytes = (nbytes + 255) // 256 * 256
        if symm_mem.get_signal_pad_size() < 16384:
            symm_mem.set_signal_pad_size(16384)  # room for two slots of 148 CTAs x 8 ranks flag words (9472 bytes)
        self.bytes = symm_mem.empty(nbytes, dtype=torch.uint8, device=device)
        self.bytes.zero_()
        self.handle = symm_mem.rendezvous(self.bytes, self.group.group_name)
        from ..ops import kernels

        self.peer_ptrs = [int(x) for x in self.handle.buffer_ptrs]
        self.signal_ptrs = [int(x) for x in self.handle.signal_pad_ptrs]
        self.peer_ptrs_c = kernels.host_ptr_array(self.peer_ptrs)
        self.signal_ptrs_c = kernels.host_ptr_array(self.signal_ptrs)
        need = 2 * 4 * kernels._lib().b200_comm_signal_words()  # two signal slots of 32-bit words
        pad = int(getattr(self.handle, "signal_pad_size", need))
        if pad < need:
            raise RuntimeError(f"symmetric-memory signal pad too small: {pad} < {need} bytes")
        try:
            self.multicast_ptr = int(self.handle.multicast_ptr)
        except Exception:  # pragma: no cover
            self.multicast_ptr = 0
        self.nbytes = nbytes
        torch.cuda.synchronize(device)
        dist.barrier(self.group)

    def view(self, byte_off: int, numel: int, dtype: torch.dtype) -> torch.Tensor:
        es = torch.empty((), dtype=dtype).element_size()
        return self.bytes.narrow(0, byte_off, numel * es).view(dtype)


class SymmReducer:
    kind = "symm"

    def __init__(self, items, group, per_parameter: bool, dtypes, devices):
        self.group = group if group is not None else dist.group.WORLD
        self.world_size = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        self.device = torch.device(devices[0])
        self.per_parameter = per_parameter
        numels = [p.numel() for p in items] if per_parameter else list(items)
        self._dtypes = list(dtypes)
        self._offs: list[int] = []
        off = 0
        for n, dt in zip(numels, dtypes):
            es = torch.empty((), dtype=dt).element_size()
            self._offs.append(off)
            off += (n * es + 255) // 256 * 256
        self._numels = numels
        self.arena = SymmArena(max(off, 256), self.device, self.group)
        self._bufs = [self.arena.view(o, n, dt) for o, n, dt in zip(self._offs, numels, dtypes)]
        # per-parameter mode: each parameter's slot is (its own flat view, offset 0)
        self.slots = [(b, 0) for b in self._bufs] if per_parameter else None
        self.stream = torch.cuda.Stream(device=self.device)
        self._seq = 0

    def buffer(self, bi: int) -> torch.Tensor:
        return self._bufs[bi]

    def _launch(self, bi: int):
        from ..ops import kernels

        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            kernels.allreduce_avg(self.arena, self._offs[bi], self._numels[bi], self._dtypes[bi],
                                  one_shot_max=_ONE_SHOT_MAX_BYTES, blocks=_COMM_BLOCKS, algo=_ALGO)
        return bi

    def reduce_bucket_async(self, bi: int):
        return self._launch(bi)

    def reduce_param_async(self, idx: int, param: torch.Tensor):
        buf = self._bufs[idx]
        g = param.grad
        if g.data_ptr() != buf.data_ptr():
            buf.view_as(g).copy_(g)
            param.grad = buf.view_as(g)
        return self._launch(idx)

    def stream_after(self, handle):
        return self.stream  # reductions are stream-ordered on the reducer's own stream

    def wait(self, handles) -> None:
        if not handles:
            return
        ev = torch.cuda.Event()
        ev.record(self.stream)
        torch.cuda.current_stream(self.device).wait_event(ev)


@torch.no_grad()
def broadcast_tensors(tensors, src: int, group=None) -> None:
    """Make ``tensors`` (same device, any dtypes) equal to rank ``src``'s copies with ONE NVLink kernel over a packed
    symmetric staging buffer (the reference issues one NCCL broadcast per parameter): a multicast store through the
    switch when the arena has an NVLS mapping (the source sends the bytes once), else every other rank pulls."""
    from ..ops import kernels

    if not tensors:
        return
    dev = tensors[0].device
    offs, off = [], 0
    for t in tensors:
        offs.append(off)
        off += (t.numel() * t.element_size() + 255) // 256 * 256
    arena = SymmArena(off, dev, group)
    views = [arena.bytes.narrow(0, o, t.numel() * t.element_size()).view(t.dtype).view(t.shape) for o, t in zip(offs, tensors)]
    if arena.rank == src:
        for v, t in zip(views, tensors):
            v.copy_(t)
    if arena.multicast_ptr and os.environ.get("B200SYS_BCAST", "mc") != "pull":
        kernels.bcast_mc(arena, 0, off, src)
    else:
        kernels.bcast_pull(arena, 0, off, src)
    if arena.rank != src:
        for v, t in zip(views, tensors):
            t.copy_(v)
    torch.cuda.synchronize(dev)
    dist.barrier(arena.group)
