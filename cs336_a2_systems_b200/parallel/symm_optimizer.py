"""Fused owner-AdamW + parameter distribution over NVLink for ``ShardedOptimizer``.

The reference's sharded step is "local optimizer.step, then one blocking broadcast per parameter"
(``ddp_bucketed_overlapped_sharded.py:335-340``).  Here, when the wrapped optimizer is an AdamW we
know the arithmetic of (this package's course-semantics ``AdamW`` or ``torch.optim.AdamW``) and the
NVLink path is legal, every rank's parameters are moved into a *symmetric arena* (identical offsets
on all ranks, peer-mapped), the bf16 GEMM shadows into a second one, and a single multi-tensor kernel
per rank updates m, v, p of the parameters it owns and stores the new fp32 value and its bf16 shadow
straight into all replicas.  Two flag barriers bracket the step (no rank writes while another still
reads old values; nobody runs ahead before all stores landed).  Optimizer state exists only on the owner.
"""

from __future__ import annotations

import torch
import torch.distributed as dist

from ..ops import kernels, native
from . import symm

_ALIGN = 256


def _mode_of(optimizer_cls) -> int | None:
    from ..optim import AdamW as CourseAdamW

    if optimizer_cls is CourseAdamW or issubclass(optimizer_cls, CourseAdamW):
        return 0
    if optimizer_cls is torch.optim.AdamW:
        return 1
    return None


def try_make(sharded, module=None):
    mode = _mode_of(sharded.optimizer_cls)
    if mode is None or sharded.world_size == 1:
        return None
    params = []
    seen = set()
    for g in sharded.param_groups:
        for p in g["params"]:
            if id(p) not in seen:
                seen.add(id(p))
                params.append(p)
    if not params:
        return None
    group = sharded.process_group
    ok, _why = symm.is_legal([p.device for p in params], [p.dtype for p in params], group)
    if not ok or any(p.dtype != torch.float32 for p in params):
        return None
    return SymmShardedAdamW(sharded, params, mode)


class SymmShardedAdamW:
    SLOT = None  # signal slot (in 32-bit words) used by the optimizer's barriers; set in __init__

    def __init__(self, sharded, params, mode: int):
        self.sharded = sharded
        self.mode = mode
        self.params = params
        dev = params[0].device
        group = sharded.process_group
        # ---- move every parameter into the symmetric arena (same order => same offsets on all ranks)
        self.offsets = []
        off = 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() * 4 + _ALIGN - 1) // _ALIGN * _ALIGN
        self.arena = symm.SymmArena(off, dev, group)
        self.shadow_arena = symm.SymmArena(off // 2, dev, group)
        self.SLOT = kernels._lib().b200_comm_signal_words()  # second slot; the first belongs to the DDP reducer
        self._off_of = {}
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                view = self.arena.view(o, p.numel(), torch.float32).view(p.shape)
                view.copy_(p.data)
                p.data = view
                self._off_of[id(p)] = o
        native.drop_shadows(params)
        native.set_shadow_provider(self._provide_shadow)
        self._plans = {}
        torch.cuda.synchronize(dev)
        dist.barrier(group)

    # joint bf16 shadow for a group of weights that sit back to back in the arena: a view of the shadow arena
    def _provide_shadow(self, weights):
        offs = [self._off_of.get(id(w)) for w in weights]
        if any(o is None for o in offs):
            return None
        for w, o, nxt in zip(weights, offs, offs[1:] + [None]):
            if nxt is not None and o + w.numel() * 4 != nxt:
                return None  # padding gap: not contiguous
        rows = sum(w.shape[0] for w in weights)
        K = weights[0].shape[1]
        return self.shadow_arena.view(offs[0] // 2, rows * K, torch.bfloat16).view(rows, K)

    def _state_for(self, p):
        inner = self.sharded.optimizer
        st = inner.state[p]
        if self.mode == 0:
            if "m" not in st:
                st["m"], st["v"], st["t"] = torch.zeros_like(p), torch.zeros_like(p), 1
            return st["m"], st["v"], st["t"]
        if "exp_avg" not in st:
            st["step"] = torch.tensor(0.0)
            st["exp_avg"], st["exp_avg_sq"] = torch.zeros_like(p), torch.zeros_like(p)
        return st["exp_avg"], st["exp_avg_sq"], int(st["step"]) + 1

    @torch.no_grad()
    def step(self, sharded):
        arena = self.arena
        # nobody may still be reading the old parameter values (forward/backward of this step are done on
        # this stream; the barrier extends that guarantee to every rank)
        kernels.symm_barrier(arena, self.SLOT)
        inner = sharded.optimizer
        if inner is not None:
            sharded._sync_hyperparameters()
            for gi, g in enumerate(inner.param_groups):
                ps = [p for p in g["params"] if p.grad is not None]
                if not ps:
                    continue
                key = tuple((p.data_ptr(), p.grad.data_ptr()) for p in ps)
                plan = self._plans.get(gi)
                if plan is None or plan[0] != key:
                    ms, vs, ts = zip(*(self._state_for(p) for p in ps))
                    shadows = []
                    for p in ps:
                        sl = [s for s in native.shadow_slices_of(p) if self._in_shadow_arena(s)]
                        shadows.append(sl[0] if sl else None)
                    launch = kernels.AdamWLaunch([p.data for p in ps], [p.grad for p in ps], list(ms), list(vs),
                                                 shadows, step0=ts[0], mode=self.mode)
                    plan = (key, launch, ps)
                    self._plans[gi] = plan
                launch = plan[1]
                b1, b2 = g["betas"]
                launch.set_hyper(g["lr"], b1, b2, g["eps"], g["weight_decay"])
                kernels.adamw_multi_distribute(launch, arena, self.shadow_arena)
                for p in ps:
                    st = inner.state[p]
                    if self.mode == 0:
                        st["t"] += 1
                    else:
                        st["step"] += 1
        kernels.symm_barrier(arena, self.SLOT)  # every owner's stores have landed everywhere
        # weights whose bf16 shadow is not arena-resident are re-cast lazily on the next forward
        for p in self.params:
            if getattr(p, "_b200_shadow_slices", None):
                for ent, i in p._b200_shadow_slices:
                    if not self._in_shadow_arena(ent.tensor):
                        ent.versions[i] = None

    def _in_shadow_arena(self, t: torch.Tensor) -> bool:
        lo = self.shadow_arena.bytes.data_ptr()
        return lo <= t.data_ptr() < lo + self.shadow_arena.nbytes
