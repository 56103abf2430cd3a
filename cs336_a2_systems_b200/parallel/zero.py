"""Bucket-fused ZeRO-1 step over NVLink: gradient reduce-scatter + sharded AdamW + parameter all-gather in ONE
kernel per DDP bucket (``csrc/comm_adamw.cu: zero_adamw_kernel``).

What the reference does per step with ``DDP_Bucketed`` + ``ShardedStateOptimizer``
(``ddp_bucketed_overlapped_sharded.py:289-318,335-340``): flatten + NCCL all-reduce of every bucket, unflatten and
divide, the owner's AdamW over whole parameters, then one blocking NCCL broadcast per parameter tensor.

Here the gradient buckets, the parameters and their bf16 GEMM shadows live in three symmetric arenas with the SAME
element layout.  When a bucket's last gradient has been produced, one kernel on the communication stream lets
rank r (a) average slice r of the bucket straight out of every rank's gradient arena, (b) apply AdamW with its
slice of m and v — optimizer state and optimizer HBM traffic are 1/world_size per rank — and (c) store the new
fp32 parameters and bf16 shadows into every rank's arenas.  ``step()`` only joins the stream.

Consequences: ``param.grad`` keeps the LOCAL (unreduced) gradient after the step, and optimizer state is sharded
by bucket slice, not by parameter (``consolidated_state_dict`` reassembles per-parameter ``m``/``v``).
"""

from __future__ import annotations

import ctypes
import os

import torch
import torch.distributed as dist

from .. import _native
from ..ops import kernels, native
from . import symm


def slice_floats(numel: int, world_size: int) -> int:
    """fp32 elements of one rank's slice of a bucket of ``numel`` elements: the kernel splits the bucket into
    32-byte vectors and gives each rank ``ceil(nvec / world_size)`` consecutive ones."""
    nvec = (numel * 4 + 31) // 32
    return (nvec + world_size - 1) // world_size * 8


def split_bucket_state(full: torch.Tensor, shapes, offsets):
    """Per-parameter views of a bucket-layout flat tensor (``offsets`` in elements, as in ``DDPBucketed``)."""
    out = []
    for shape, off in zip(shapes, offsets):
        n = 1
        for d in shape:
            n *= d
        out.append(full[off:off + n].view(shape))
    return out


def assemble_bucket_state(tensors, offsets, total: int, device=None) -> torch.Tensor:
    """Inverse of :func:`split_bucket_state`: a zero-padded flat tensor of ``total`` elements holding
    ``tensors[i]`` at ``offsets[i]`` (None entries stay zero)."""
    full = torch.zeros(total, dtype=torch.float32, device=device)
    for t, off in zip(tensors, offsets):
        if t is not None:
            full[off:off + t.numel()].copy_(t.reshape(-1))
    return full


class BucketFusedAdamW:
    def __init__(self, sharded, ddp, mode: int):
        red = ddp._reducer
        self.sharded, self.ddp, self.mode = sharded, ddp, mode
        self.garena = red.arena
        self.rank, self.ws = self.garena.rank, self.garena.world_size
        dev = red.device
        group = red.group
        self.device = dev
        nbytes = self.garena.nbytes
        self.parena = symm.SymmArena(nbytes, dev, group)
        self.sarena = symm.SymmArena(nbytes // 2, dev, group)
        self._off_of: dict[int, int] = {}
        self._buckets = []
        with torch.no_grad():
            for bi, b in enumerate(ddp.buckets):
                assert b.dtype == torch.float32, "bucket-fused AdamW needs fp32 parameters"
                boff = red._offs[bi]
                for p, off in zip(b.params, b.offsets):
                    byte = boff + off * 4
                    view = self.parena.view(byte, p.numel(), torch.float32).view(p.shape)
                    view.copy_(p.data)
                    p.data = view
                    self._off_of[id(p)] = byte
                per8 = slice_floats(b.numel, self.ws)
                m = torch.zeros(per8, dtype=torch.float32, device=dev)
                v = torch.zeros(per8, dtype=torch.float32, device=dev)
                self._buckets.append((boff, b.numel, m, v, per8))
        self._params = [p for b in ddp.buckets for p in b.params]
        native.drop_shadows(self._params)
        native.set_shadow_provider(self._provide_shadow)
        probe = torch.nn.Parameter(torch.zeros(1))
        self._defaults = dict(sharded.optimizer_cls([probe], **sharded.optimizer_kwargs).defaults)
        # measured per-variant on B200 (profiles/zero_kernel_n*.json): with 2 ranks plain P2P moves the fewest
        # bytes and the bf16 shadows ride along; from 3 ranks up the shadows are re-cast locally from the fp32
        # parameters that arrived (fewer NVLink bytes) and the gradient slice can be reduced in the switch
        mcs = [self.garena.multicast_ptr, self.parena.multicast_ptr, self.sarena.multicast_ptr]
        have_mc = all(mcs) and os.environ.get("B200SYS_MULTIMEM", "1") != "0"
        self._mc = kernels.host_ptr_array(mcs) if have_mc else None
        v = os.environ.get("B200SYS_ZERO_VARIANT", "auto")
        if v == "auto":
            # the all-NVLS variant (3) is the fastest stand-alone at 8 ranks (2.19 vs 2.82 ms/GB) but has only been
            # exercised in the micro-benchmark; the P2P variant is the one validated end to end at 2/4/8 ranks
            self.variant = 0
        else:
            self.variant = int(v) if have_mc else 0
        sh = os.environ.get("B200SYS_ZERO_SHADOW", "auto")
        self.wire_shadows = (self.ws <= 2) if sh == "auto" else (sh == "wire")
        self.hyper = torch.zeros(8, dtype=torch.float32, device=dev)
        self._hyper_host = torch.zeros(16, 8, dtype=torch.float32).pin_memory()
        self._hyper_slot = 0
        self._hyper_vals = None
        self.step_dev = torch.ones(1, dtype=torch.int32, device=dev)
        self.t = 1
        self._launched = False
        ddp.set_bucket_reduce_override(self._reduce_bucket)
        torch.cuda.synchronize(dev)
        dist.barrier(group)

    # bf16 shadow of a group of weights that are adjacent in the parameter arena = a view of the shadow arena
    def _provide_shadow(self, weights):
        offs = [self._off_of.get(id(w)) for w in weights]
        if any(o is None for o in offs):
            return None
        for w, o, nxt in zip(weights, offs, offs[1:] + [None]):
            if nxt is not None and o + w.numel() * 4 != nxt:
                return None
        rows = sum(w.shape[0] for w in weights)
        return self.sarena.view(offs[0] // 2, rows * weights[0].shape[1], torch.bfloat16).view(rows, weights[0].shape[1])

    def _set_hyper(self):
        g = dict(self._defaults)
        g.update({k: v for k, v in self.sharded.param_groups[0].items() if k != "params"})  # LR schedules write here
        b1, b2 = g["betas"]
        vals = (float(g["lr"]), float(b1), float(b2), float(g["eps"]), float(g["weight_decay"]), 1.0)
        if vals != self._hyper_vals:
            self._hyper_vals = vals
            row = self._hyper_host[self._hyper_slot]
            self._hyper_slot = (self._hyper_slot + 1) % self._hyper_host.shape[0]
            row[:6] = torch.tensor(vals)
            self.hyper.copy_(row, non_blocking=True)

    def refresh_device_hyper(self) -> None:
        """See ``AdamW.refresh_device_hyper``: called by the trainer before a CUDA-graph replay."""
        self._set_hyper()

    def _reduce_bucket(self, bi: int):
        """Called by DDPBucketed in place of its all-reduce when bucket ``bi`` is complete."""
        red = self.ddp._reducer
        cur = torch.cuda.current_stream(self.device)
        if not self._launched:
            self._set_hyper()  # on the compute stream, ordered before the event below
        ev = torch.cuda.Event()
        ev.record(cur)
        red.stream.wait_event(ev)
        boff, numel, m, v, _per = self._buckets[bi]
        with torch.cuda.stream(red.stream):
            _native.check(_native.lib().b200_zero_adamw_bucket(
                self.garena.peer_ptrs_c, self.parena.peer_ptrs_c, self.sarena.peer_ptrs_c if self.wire_shadows else None,
                self.garena.signal_ptrs_c, self._mc, ctypes.c_longlong(boff), ctypes.c_longlong(numel), _native.p(m),
                _native.p(v), _native.p(self.hyper), _native.p(self.step_dev), ctypes.c_int(self.mode),
                ctypes.c_int(self.rank), ctypes.c_int(self.ws), ctypes.c_int(0), ctypes.c_int(0),
                ctypes.c_int(self.variant), ctypes.c_void_p(red.stream.cuda_stream)), "zero_adamw_bucket")
            if not self.wire_shadows:
                # every rank's slice of the new fp32 parameters has landed (trailing barrier of the kernel)
                kernels.cast_bf16(self.parena.view(boff, numel, torch.float32),
                                  self.sarena.view(boff // 2, numel, torch.bfloat16))
            kernels._count()
        self._launched = True
        return bi

    @torch.no_grad()
    def step(self, sharded):
        red = self.ddp._reducer
        if self._launched:
            with torch.cuda.stream(red.stream):
                _native.check(_native.lib().b200_incr_step(_native.p(self.step_dev),
                                                           ctypes.c_void_p(red.stream.cuda_stream)), "incr_step")
            ev = torch.cuda.Event()
            ev.record(red.stream)
            torch.cuda.current_stream(self.device).wait_event(ev)
            self.t += 1
            self._launched = False
            kernels.bump_param_epoch()
            # shadows that could not live in the arena (a fused group straddling a bucket edge) are re-cast lazily
            lo = self.sarena.bytes.data_ptr()
            hi = lo + self.sarena.nbytes
            for p in self._params:
                for ent, i in getattr(p, "_b200_shadow_slices", ()):
                    if not (lo <= ent.tensor.data_ptr() < hi):
                        ent.versions[i] = None

    # ------------------------------------------------------------------ state for checkpoints
    def _param_index(self):
        index = {}
        i = 0
        for g in self.sharded.param_groups:
            for p in g["params"]:
                index[id(p)] = i
                i += 1
        return index

    def full_state(self):
        """``{param_index: {"m", "v", "t"}}`` reassembled from the bucket slices (collective)."""
        out = {}
        index = self._param_index()
        t = int(self.step_dev.item())  # device-resident: stays right under CUDA-graph replay
        for bi, b in enumerate(self.ddp.buckets):
            _boff, _numel, m, v, per8 = self._buckets[bi]
            full_m = torch.empty(per8 * self.ws, device=self.device)
            full_v = torch.empty(per8 * self.ws, device=self.device)
            dist.all_gather_into_tensor(full_m, m, group=self.garena.group)
            dist.all_gather_into_tensor(full_v, v, group=self.garena.group)
            shapes = [p.shape for p in b.params]
            for p, pm, pv in zip(b.params, split_bucket_state(full_m, shapes, b.offsets),
                                 split_bucket_state(full_v, shapes, b.offsets)):
                if id(p) in index:
                    out[index[id(p)]] = {"m": pm.cpu(), "v": pv.cpu(), "t": t}
        return out

    @torch.no_grad()
    def load_full_state(self, state: dict) -> None:
        """Inverse of :meth:`full_state`: every rank keeps its slice of each bucket's m and v (no communication;
        every rank reads the same consolidated dict)."""
        index = self._param_index()
        t = None
        for bi, b in enumerate(self.ddp.buckets):
            _boff, _numel, m, v, per8 = self._buckets[bi]
            sts = [state.get(index.get(id(p), -1)) or state.get(str(index.get(id(p), -1))) for p in b.params]
            for key, dst in (("m", m), ("v", v)):
                full = assemble_bucket_state([None if st is None else st[key].to(torch.float32) for st in sts],
                                             b.offsets, per8 * self.ws)
                dst.copy_(full[self.rank * per8:(self.rank + 1) * per8])
            for st in sts:
                if st is not None and "t" in st:
                    t = int(st["t"])
        if t is not None:
            self.step_dev.fill_(t)
            self.t = t
