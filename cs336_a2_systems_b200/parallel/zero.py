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

Consequences: ``param.grad`` keeps the LOCAL (unreduced) gradient after the step (``ddp.grads_are_local`` is set),
and optimizer state is sharded by bucket slice, not by parameter (``consolidated_state_dict`` reassembles
per-parameter ``m``/``v``).  The path is opt-in (``ShardedOptimizer.enable_symmetric(ddp_model)``); without that call
the wrapper keeps the reference's semantics (averaged ``.grad`` on every rank, whole-parameter ownership).

Deferred fp32 gather (``defer_fp32_gather=True``): the bucket kernels put only the bf16 GEMM shadows on the wire — the
only copies the next forward's GEMMs read — plus the fp32 values of the parameters forward reads in fp32 (RMSNorm gains,
the embedding table).  The fp32 replicas of the remaining (Linear) weights are refreshed by ONE gather kernel that runs
on the communication stream under the next forward (launched by the DDP wrapper's forward pre-hook), or on demand by
``sync_params()`` / ``state_dict()``.  Per-rank NVLink traffic inside the backward window drops from 6 to 2 bytes per
parameter; the cost is that between ``optimizer.step()`` and the next forward (or ``sync_params()``) a rank's fp32
copies of Linear weights it does not own are one step old.
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


MAX_EAGER_RANGES = 4  # == kZEager in csrc/comm_adamw.cu (the C side rejects more)


def eager_ranges(params, offsets, is_eager, align: int = 8):
    """Element ranges ``[lo, hi)`` of a bucket covering the parameters for which ``is_eager(p)`` holds, with
    neighbours merged (slots are ``align``-element aligned, so two adjacent slots touch after rounding up)."""
    out: list[list[int]] = []
    for p, off in zip(params, offsets):
        if not is_eager(p):
            continue
        lo, hi = off, off + (p.numel() + align - 1) // align * align
        if out and out[-1][1] == lo:
            out[-1][1] = hi
        else:
            out.append([lo, hi])
    return [(a, b) for a, b in out]


class BucketFusedAdamW:
    def __init__(self, sharded, ddp, mode: int, defer_fp32_gather: bool | None = None, variant: int | None = None):
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
        v = os.environ.get("B200SYS_ZERO_VARIANT", "auto") if variant is None else str(int(variant))
        if v == "auto":
            # Peer loads/stores at every world size.  From 3 ranks up the all-NVLS variant (3: the switch reduces the
            # gradient slice and replicates the stores) is faster when it runs — 2.19 vs 2.82 ms/GB stand-alone, 47.7 vs
            # 52.1 ms/step end to end at 8 ranks (profiles/r2) — but end-to-end runs with the NVLS variants stopped making
            # progress at 8 ranks in one call of round 1 and in one call of round 2 (three runs in a row, no barrier
            # time-out record; cause not identified), while the peer-to-peer variant has never stalled.  Opt in with
            # variant=3 / B200SYS_ZERO_VARIANT=3.
            self.variant = 0
        else:
            self.variant = int(v) if have_mc else 0
        #: CTAs per bucket kernel (0 = one per SM): fewer CTAs leave more of each SM's issue slots and memory queues to
        #: the backward GEMMs the kernel overlaps with
        self.blocks = int(os.environ.get("B200SYS_ZERO_BLOCKS", "0"))
        if defer_fp32_gather is None:
            defer_fp32_gather = os.environ.get("B200SYS_ZERO_DEFER", "0") == "1"
        #: "lazy": the fp32 replicas of Linear weights a rank does not own are refreshed only by sync_params() /
        #: state_dict() (fp32 master weights live sharded with the optimizer state, as in ZeRO with bf16 parameters)
        self.lazy = defer_fp32_gather == "lazy"
        self.defer = bool(defer_fp32_gather)
        sh = os.environ.get("B200SYS_ZERO_SHADOW", "auto")
        self.wire_shadows = True if self.defer else ((self.ws <= 2) if sh == "auto" else (sh == "wire"))
        # parameters that forward reads in fp32: everything that is not a Linear weight, plus (added by the shadow
        # provider when it has to decline a group) Linear weights whose bf16 shadow cannot live in the shadow arena
        from ..models.layers import Linear as _Linear

        shadowed = {id(m.weight) for m in ddp.module.modules() if isinstance(m, _Linear)}
        self._eager_ids = {id(q) for b in ddp.buckets for q in b.params if id(q) not in shadowed}
        self._rebuild_eager()
        self._gather_pending = False
        self._gather_event = None
        if self.defer:
            if not self.lazy:
                ddp.register_forward_pre_hook(lambda _m, _a: self.launch_param_gather())
            ddp.module.register_state_dict_pre_hook(lambda *_a, **_k: self.sync_params())
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

    def _rebuild_eager(self) -> None:
        """Per-bucket fp32-eager ranges + the table of buckets whose fp32 gather is deferred."""
        red = self.ddp._reducer
        self._eager, self._defer_bucket = [], []
        for b in self.ddp.buckets:
            rs = eager_ranges(b.params, b.offsets, lambda q: id(q) in self._eager_ids) if self.defer else []
            ok = self.defer and len(rs) <= MAX_EAGER_RANGES  # too fragmented: this bucket distributes fp32 at once
            self._defer_bucket.append(ok)
            rs = rs if ok else []
            lo = (ctypes.c_longlong * MAX_EAGER_RANGES)(*([r[0] for r in rs] + [0] * (MAX_EAGER_RANGES - len(rs))))
            hi = (ctypes.c_longlong * MAX_EAGER_RANGES)(*([r[1] for r in rs] + [0] * (MAX_EAGER_RANGES - len(rs))))
            self._eager.append((len(rs), lo, hi))
        self._gather_table = torch.tensor([[red._offs[bi], (b.numel * 4 + 31) // 32]
                                           for bi, b in enumerate(self.ddp.buckets) if self._defer_bucket[bi]],
                                          dtype=torch.int64).reshape(-1, 2).to(self.device)

    def _decline_shadow(self, weights):
        """The group's bf16 shadow will live outside the arena and be re-cast from the fp32 parameters: with a
        deferred gather those must then be distributed eagerly."""
        if self.defer and not all(id(w) in self._eager_ids for w in weights):
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("a new weight group appeared during CUDA-graph capture; run one eager step first")
            self._eager_ids.update(id(w) for w in weights)
            self._rebuild_eager()
        return None

    # bf16 shadow of a group of weights that are adjacent in the parameter arena = a view of the shadow arena
    def _provide_shadow(self, weights):
        offs = [self._off_of.get(id(w)) for w in weights]
        if any(o is None for o in offs):
            return self._decline_shadow(weights)
        for w, o, nxt in zip(weights, offs, offs[1:] + [None]):
            if nxt is not None and o + w.numel() * 4 != nxt:
                return self._decline_shadow(weights)
        rows = sum(w.shape[0] for w in weights)
        return self.sarena.view(offs[0] // 2, rows * weights[0].shape[1], torch.bfloat16).view(rows, weights[0].shape[1])

    def _set_hyper(self):
        g = dict(self._defaults)
        g.update({k: v for k, v in self.sharded.param_groups[0].items() if k != "params"})  # LR schedules write here
        b1, b2 = g["betas"]
        vals = (float(g["lr"]), float(b1), float(b2), float(g["eps"]), float(g["weight_decay"]), 1.0)
        if vals != self._hyper_vals:
            if torch.cuda.is_current_stream_capturing():
                return  # never record the staging copy into a graph (see kernels.AdamWLaunch.set_hyper)
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
                ctypes.c_int(self.rank), ctypes.c_int(self.ws), ctypes.c_int(self.blocks), ctypes.c_int(0),
                ctypes.c_int(self.variant), ctypes.c_int(int(self._defer_bucket[bi])), ctypes.c_int(self._eager[bi][0]),
                self._eager[bi][1], self._eager[bi][2], ctypes.c_void_p(red.stream.cuda_stream)), "zero_adamw_bucket")
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
            self._gather_pending = self.defer and self._gather_table.shape[0] > 0
            kernels.bump_param_epoch()
            # shadows that could not live in the arena (a fused group straddling a bucket edge) are re-cast lazily
            lo = self.sarena.bytes.data_ptr()
            hi = lo + self.sarena.nbytes
            for p in self._params:
                for ent, i in getattr(p, "_b200_shadow_slices", ()):
                    if not (lo <= ent.tensor.data_ptr() < hi):
                        ent.versions[i] = None

    # ------------------------------------------------------------------ deferred fp32 gather
    def launch_param_gather(self) -> None:
        """Enqueue the deferred fp32 all-gather of the last step on the communication stream (no-op when nothing is
        pending).  Does NOT make the current stream wait: forward reads the bf16 shadows and the eager fp32 ranges,
        both already distributed by the bucket kernels."""
        if not self._gather_pending:
            return
        red = self.ddp._reducer
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(self.device))
        red.stream.wait_event(ev)  # (also what pulls the communication stream into a CUDA-graph capture)
        mc = self.parena.multicast_ptr if (self.variant & 2) and self._mc is not None else 0
        with torch.cuda.stream(red.stream):
            _native.check(_native.lib().b200_zero_param_gather(
                self.parena.peer_ptrs_c, self.garena.signal_ptrs_c, ctypes.c_ulonglong(mc),
                _native.p(self._gather_table), ctypes.c_int(self._gather_table.shape[0]), ctypes.c_int(self.rank),
                ctypes.c_int(self.ws), ctypes.c_int(0), ctypes.c_int(0), ctypes.c_void_p(red.stream.cuda_stream)),
                "zero_param_gather")
            kernels._count()
        self._gather_event = torch.cuda.Event()
        self._gather_event.record(red.stream)
        self._gather_pending = False

    def sync_params(self) -> None:
        """Make every fp32 parameter replica current on this rank, ordered before later work on the current stream
        (collective: every rank must call it; ``state_dict()`` of the wrapped module does so itself)."""
        self.launch_param_gather()
        if self._gather_event is not None:
            torch.cuda.current_stream(self.device).wait_event(self._gather_event)
            self._gather_event = None

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
