"""AdamW with the course library's arithmetic and state layout, plus a fused sm_100a path.

Semantics (``cs336_basics/optimizer.py:30-86``): per-parameter state ``{"m", "v", "t"}`` with
``t`` starting at 1; ``α_t = lr·√(1-β₂ᵗ)/(1-β₁ᵗ)``; ``p -= α_t·m/(√v+eps)`` and then
``p -= lr·wd·p`` (decay AFTER the update — not ``torch.optim.AdamW``'s arithmetic).

CUDA fp32 parameters take ONE multi-tensor kernel launch for the whole model
(``csrc/adamw.cu``) instead of ~12 ATen launches per tensor: it reads a device-resident
descriptor table, keeps the step counter and the hyper-parameters in device memory (so the
step can live inside a CUDA graph and the LR schedule is a 4-byte async copy) and, in the
same pass, refreshes the bf16 weight shadows the tcgen05 GEMMs read — the autocast
weight-cast kernels disappear from the step.
"""

from __future__ import annotations

import math
from collections.abc import Callable, Iterable

import torch

from .. import _native


def get_cosine_lr(it: int, max_learning_rate: float, min_learning_rate: float, warmup_iters: int,
                  cosine_cycle_iters: int) -> float:
    """Linear warm-up then cosine decay to ``min_learning_rate`` (optimizer.py:9-27)."""
    if it < warmup_iters:
        return max_learning_rate * it / warmup_iters
    if it > cosine_cycle_iters:
        return min_learning_rate
    ratio = (it - warmup_iters) / (cosine_cycle_iters - warmup_iters)
    coeff = 0.5 * (1.0 + math.cos(math.pi * ratio))
    return min_learning_rate + coeff * (max_learning_rate - min_learning_rate)


class AdamW(torch.optim.Optimizer):
    def __init__(self, params: Iterable, lr: float = 1e-3, betas: tuple[float, float] = (0.9, 0.999),
                 eps: float = 1e-8, weight_decay: float = 0.01, fused: bool | None = None):
        if lr < 0.0:
            raise ValueError(f"Invalid learning rate: {lr}")
        if eps < 0.0:
            raise ValueError(f"Invalid epsilon value: {eps}")
        if not 0.0 <= betas[0] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 0: {betas[0]}")
        if not 0.0 <= betas[1] < 1.0:
            raise ValueError(f"Invalid beta parameter at index 1: {betas[1]}")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._fused = fused
        self._plans: dict[object, object] = {}
        self._done: set[int] = set()  # ids of params already stepped this iteration by step_subset()
        self._bwd_groups: dict[int, "_BackwardFusedGroup"] = {}  # see fuse_into_backward()

    # ------------------------------------------------------------------ reference arithmetic
    @staticmethod
    def _step_param_torch(p, state, lr, b1, b2, eps, wd):
        g = p.grad.data
        if g.is_sparse:
            raise RuntimeError("AdamW does not support sparse gradients")
        t = state.get("t", 1)
        m = state.get("m")
        v = state.get("v")
        if m is None:
            m = torch.zeros_like(g)
            v = torch.zeros_like(g)
        m = b1 * m + (1 - b1) * g
        v = b2 * v + (1 - b2) * torch.square(g)
        alpha_t = lr * (math.sqrt(1 - b2 ** t) / (1 - b1 ** t))
        # in place on the parameter itself (callers hold no_grad): unlike ``p.data -= ...`` this bumps
        # ``p._version``, which is what tells the bf16 weight-shadow cache (ops/native.py) to re-cast
        p.sub_(alpha_t * m / (torch.sqrt(v) + eps))
        p.sub_(lr * wd * p)
        state["m"], state["v"], state["t"] = m, v, t + 1
        if p.is_cuda:
            from ..ops import kernels
            kernels.bump_param_epoch()

    def _fusable(self, p) -> bool:
        if self._fused is False or _native.force_torch():
            return False
        return p.is_cuda and p.dtype == torch.float32 and p.grad is not None and \
            p.grad.dtype == torch.float32 and not p.grad.is_sparse and p.is_contiguous() and \
            p.grad.is_contiguous()

    @torch.no_grad()
    def step(self, closure: Callable | None = None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            lr, (b1, b2), eps, wd = group["lr"], group["betas"], group["eps"], group["weight_decay"]
            fused_ps = []
            for p in group["params"]:
                if p.grad is None or id(p) in self._done:
                    continue
                if self._fusable(p):
                    fused_ps.append(p)
                else:
                    self._step_param_torch(p, self.state[p], lr, b1, b2, eps, wd)
            if fused_ps:
                self._step_fused(gi, fused_ps, lr, b1, b2, eps, wd)
        for grp in self._bwd_groups.values():
            grp.finish_step()
        self._done.clear()
        return loss

    # ------------------------------------------------------------------ optimizer inside the wgrad GEMM epilogue
    def fuse_into_backward(self, module: torch.nn.Module | None = None, enable: bool = True) -> int:
        """Let the weight-gradient GEMMs of the native ``Linear`` layers apply this optimizer in their epilogue
        (``csrc/gemm_epilogue.cuh: EPI_ADAMW``): the fp32 dW tile in the accumulator is combined with p, m, v of the
        same elements and p, m, v and the bf16 GEMM shadow are written back — the gradient of those weights never
        reaches HBM (no 4 B/param write in the GEMM, no 4 B/param read in the optimizer) and their update overlaps
        the tensor-core work of backward.  ``step()`` then only advances the step counter for them (and updates
        everything else: norm gains, embeddings, …).

        Same arithmetic and state (``m``, ``v``, ``t``) as ``step()``.  Restrictions: ``param.grad`` stays ``None``
        for fused weights (so no gradient clipping / accumulation / inspection — the caller must not need them),
        every fused weight must take part in every backward, and hyper-parameters are read at backward time.
        Returns the number of parameters registered; ``enable=False`` undoes it."""
        from ..models.layers import Linear

        for grp in self._bwd_groups.values():
            grp.release()
        self._bwd_groups = {}
        if not enable or self._fused is False or _native.force_torch():
            return 0
        wanted = None
        if module is not None:
            wanted = {id(m.weight) for m in module.modules() if isinstance(m, Linear)}
        n = 0
        for gi, g in enumerate(self.param_groups):
            ps = [p for p in g["params"] if p.is_cuda and p.dtype == torch.float32 and p.dim() == 2 and
                  p.is_contiguous() and p.requires_grad and (wanted is None or id(p) in wanted)]
            if ps:
                self._bwd_groups[gi] = _BackwardFusedGroup(self, gi, ps)
                n += len(ps)
        return n


    @torch.no_grad()
    def step_subset(self, key, params, max_blocks: int = 0) -> None:
        """Update only ``params`` now (one fused launch on the current stream); the next :meth:`step`
        skips them.  This is what lets a trainer run the optimizer for a layer on a side stream as soon
        as that layer's gradients are final, overlapped with the rest of backward ("optimizer in
        backward"); the arithmetic is exactly that of :meth:`step`."""
        group_of = getattr(self, "_group_of", None)
        if group_of is None:
            group_of = {id(p): gi for gi, g in enumerate(self.param_groups) for p in g["params"]}
            self._group_of = group_of
        by_group: dict[int, list] = {}
        for p in params:
            if p.grad is not None and id(p) not in self._done:
                by_group.setdefault(group_of[id(p)], []).append(p)
        for gi, ps in by_group.items():
            g = self.param_groups[gi]
            lr, (b1, b2), eps, wd = g["lr"], g["betas"], g["eps"], g["weight_decay"]
            fus = [p for p in ps if self._fusable(p)]
            for p in ps:
                if not self._fusable(p):
                    self._step_param_torch(p, self.state[p], lr, b1, b2, eps, wd)
            if fus:
                self._step_fused((gi, key), fus, lr, b1, b2, eps, wd, max_blocks)
            self._done.update(id(p) for p in ps)

    # ------------------------------------------------------------------ fused sm_100a path
    def _step_fused(self, gi, ps, lr, b1, b2, eps, wd, max_blocks: int = 0):
        from ..ops import native

        for p in ps:
            st = self.state[p]
            if "m" not in st:
                st["m"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["v"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["t"] = 1
        plan = self._plans.get(gi)
        # the plan holds raw pointers: any replaced tensor (new grads, moments swapped in by load_state_dict) or a
        # step counter that no longer matches the device copy (state restored to an earlier step) rebuilds it
        key = tuple((p.data_ptr(), p.grad.data_ptr(), self.state[p]["m"].data_ptr(), self.state[p]["v"].data_ptr())
                    for p in ps)
        if plan is not None and plan.key == key and plan.host_step != self.state[ps[0]]["t"]:
            plan = None
        if plan is None or plan.key != key:
            t0 = self.state[ps[0]]["t"]
            if any(self.state[p]["t"] != t0 for p in ps):
                # mixed step counts (param added later): one plan per distinct t
                for p in ps:
                    self._step_param_torch(p, self.state[p], lr, b1, b2, eps, wd)
                return
            plan = native.AdamWPlan(ps, [p.grad for p in ps], [self.state[p]["m"] for p in ps],
                                    [self.state[p]["v"] for p in ps], key=key, step0=t0)
            self._plans[gi] = plan
        plan.set_hyper(lr, b1, b2, eps, wd)
        plan.launch(max_blocks)
        for p in ps:  # host mirror of the device step counter (exact outside CUDA-graph replay)
            self.state[p]["t"] += 1
        plan.host_step = self.state[ps[0]]["t"]

    def load_state_dict(self, state_dict) -> None:
        """Restoring state replaces the moment tensors and step counters the fused plans point at: drop the plans so
        the next step rebuilds them around the restored tensors (and a device step counter equal to the restored t)."""
        super().load_state_dict(state_dict)
        self._plans.clear()
        self._done.clear()
        for grp in self._bwd_groups.values():
            grp.reload_state()

    def __setstate__(self, state) -> None:
        super().__setstate__(state)
        self.__dict__.setdefault("_fused", None)
        self.__dict__["_plans"] = {}
        self.__dict__["_done"] = set()

    def refresh_device_hyper(self) -> None:
        """Push the current ``lr`` / ``betas`` / ``eps`` / ``weight_decay`` of every param group to the device copies the
        fused kernels read.  ``step()`` does this itself; a CUDA-graph replay does not run ``step()``, so the trainer
        calls this before each replay — that is what keeps LR schedules working under ``cuda_graph=True``.  Copies
        happen only when a value changed."""
        for k, plan in self._plans.items():
            g = self.param_groups[k if isinstance(k, int) else k[0]]
            b1, b2 = g["betas"]
            plan.set_hyper(g["lr"], b1, b2, g["eps"], g["weight_decay"])
        for grp in self._bwd_groups.values():
            grp.push_hyper()

    def sync_step_from_device(self) -> None:
        """After CUDA-graph replays the host ``t`` mirrors are stale; refresh them from the device."""
        for _key, plan in self._plans.items():
            t = plan.device_step()
            for p in plan.params:
                self.state[p]["t"] = t
            plan.host_step = t
        for grp in self._bwd_groups.values():
            grp.sync_from_device()


class _BackwardFusedGroup:
    """The weights of one param group whose AdamW update runs inside their wgrad GEMM (``AdamW.fuse_into_backward``).
    Holds the device-resident hyper-parameters / step counter the epilogues read, and marks every weight with a
    back-pointer (``_b200_bwd_opt``) that ``ops.native.LinearFn.backward`` looks for."""

    def __init__(self, opt: AdamW, gi: int, params):
        from ..ops import kernels

        self.opt, self.gi, self.params = opt, gi, list(params)
        t0 = None
        for p in self.params:
            st = opt.state[p]
            if "m" not in st:
                st["m"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["v"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                st["t"] = 1
            t0 = st["t"] if t0 is None else t0
            if st["t"] != t0:
                raise ValueError("fuse_into_backward: the weights of a param group must share one step count")
            p._b200_bwd_opt = self
        self.dh = kernels.DeviceHyper(self.params[0].device, step0=t0 or 1)
        self.touched = False
        self.push_hyper()

    def release(self) -> None:
        for p in self.params:
            if getattr(p, "_b200_bwd_opt", None) is self:
                del p._b200_bwd_opt

    def push_hyper(self) -> None:
        g = self.opt.param_groups[self.gi]
        b1, b2 = g["betas"]
        self.dh.set(g["lr"], b1, b2, g["eps"], g["weight_decay"])

    def wgrad(self, dy2, x2, weights, shadow, grad_out=None) -> None:
        """Called from ``LinearFn.backward`` in place of the plain wgrad GEMM."""
        from ..ops import kernels

        if not self.touched:  # first weight of this backward: hyper-parameters as they are NOW
            self.push_hyper()
            self.touched = True
        st = self.opt.state
        kernels.wgrad_adamw(dy2, x2, [w.data for w in weights], [st[w]["m"] for w in weights],
                            [st[w]["v"] for w in weights], shadow, self.dh, mode=0, grad_out=grad_out)

    def finish_step(self) -> None:
        from ..ops import kernels

        if not self.touched:
            return
        self.dh.advance()
        for p in self.params:
            self.opt.state[p]["t"] += 1
        self.touched = False
        kernels.bump_param_epoch()

    def reload_state(self) -> None:
        """After ``load_state_dict``: the moments are new tensors (looked up per launch, nothing cached) and the device
        step counter must equal the restored ``t``."""
        t = self.opt.state[self.params[0]].get("t", 1)
        self.dh.step.fill_(int(t))
        for p in self.params:
            p._b200_bwd_opt = self

    def sync_from_device(self) -> None:
        t = int(self.dh.step.item())
        for p in self.params:
            self.opt.state[p]["t"] = t
