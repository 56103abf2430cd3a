"""The training step: zero_grad → autocast forward → cross-entropy → backward (DDP hooks fire) →
finish gradient sync → optional clip → optimizer step.

Counterpart of the hot loops in the reference drivers
(``ddp_bucketed_overlapped_sharded.py:161-197``, ``naive_ddp.py:53-171``, ``benchmark.py:89-117``).
``Trainer.step_from_host`` is the end-to-end call a user makes: it stages the batch from pinned
host memory, runs the step and returns the loss as a Python float (one 4-byte D2H read).

With ``cuda_graph=True`` the whole step (forward, backward, reductions on their side stream,
fused AdamW) is captured once into a CUDA graph and replayed — launch-bound Python/autograd
overhead disappears; the step counter and hyper-parameters live in device memory so the replay
stays exact (see ``optim/adamw.py``).
"""

from __future__ import annotations

import contextlib
import os

import torch
from torch import Tensor

from ..ops import cross_entropy, grad_slots
from ..ops import kernels as kernels_mod
from ..ops.api import clip_gradient
from ..utils import nvtx


class GradArena:
    """Flat fp32 gradient buffer for single-process training: registers one slot per parameter so
    weight gradients are produced in place (stable addresses → CUDA-graph friendly, one flat
    buffer for clipping / the fused optimizer)."""

    def __init__(self, params, align_elems: int = 8):  # 32-byte slots: the optimizer streams 256-bit vectors
        self.params = [p for p in params if p.requires_grad]
        total = 0
        self.offsets = []
        for p in self.params:
            self.offsets.append(total)
            total += (p.numel() + align_elems - 1) // align_elems * align_elems
        dev = self.params[0].device
        self.flat = torch.zeros(total, dtype=torch.float32, device=dev)
        for p, off in zip(self.params, self.offsets):
            if p.dtype == torch.float32:
                grad_slots.register(p, self.flat, off)

    def release(self):
        grad_slots.release_all(self.params)


class _OptimizerOverlap:
    """Runs the fused AdamW for a group of parameters on a side stream as soon as their gradients are
    final, while backward continues (memory-bound optimizer under compute-bound GEMMs).

    Single process: parameters are grouped into units of ~``unit_mb`` in reverse order and a
    post-accumulate-grad hook launches a unit when its last gradient has been produced.  Under DDP the
    unit is a gradient bucket and the launch is chained behind that bucket's all-reduce.
    Requires every parameter to receive exactly one gradient per step (true for the LM)."""

    def __init__(self, model, optimizer, device, unit_mb: float, is_ddp: bool):
        self.optimizer = optimizer
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self._used = False
        self.max_blocks = 2 * torch.cuda.get_device_properties(device).multi_processor_count
        if is_ddp and hasattr(model, "set_bucket_callback"):
            model.set_bucket_callback(self._on_bucket)
            self.units = None
            return
        params = [p for p in model.parameters() if p.requires_grad]
        cap = unit_mb * 1024 * 1024
        self.units: list[list] = []
        cur, cur_b = [], 0
        for p in reversed(params):
            cur.append(p)
            cur_b += p.numel() * p.element_size()
            if cur_b >= cap:
                self.units.append(cur)
                cur, cur_b = [], 0
        if cur:
            self.units.append(cur)
        self._count = [0] * len(self.units)
        for ui, unit in enumerate(self.units):
            for p in unit:
                p.register_post_accumulate_grad_hook(self._make_hook(ui))

    def _make_hook(self, ui: int):
        def hook(param):
            self._count[ui] += 1
            if self._count[ui] == len(self.units[ui]):
                self._count[ui] = 0
                self._launch(ui, self.units[ui], torch.cuda.current_stream(self.device))
        return hook

    def _launch(self, key, params, after_stream) -> None:
        ev = torch.cuda.Event()
        ev.record(after_stream)
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            # two CTAs per SM: enough loads in flight to stream HBM, small enough to sit beside a GEMM CTA
            self.optimizer.step_subset(key, params, max_blocks=self.max_blocks)
        self._used = True

    def _on_bucket(self, bucket_index: int, params, reduce_stream) -> None:
        self._launch(("bucket", bucket_index), params, reduce_stream)

    def join(self) -> None:
        if self._used:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            self._used = False


class Trainer:
    def __init__(self, model: torch.nn.Module, optimizer: torch.optim.Optimizer, *, autocast: bool = True,
                 max_grad_norm: float | None = None, cuda_graph: bool = False, grad_arena: bool | None = None,
                 overlap_optimizer: bool = False, overlap_unit_mb: float = 128.0, graph_warmup_steps: int = 3,
                 optimizer_in_backward: bool | None = None):
        self.model = model
        self.optimizer = optimizer
        self.autocast = autocast
        self.max_grad_norm = max_grad_norm
        self.cuda_graph = cuda_graph
        self.device = next(model.parameters()).device
        self._is_ddp = hasattr(model, "finish_gradient_synchronization")
        red = getattr(model, "_reducer", None)
        if self.cuda_graph and red is not None and getattr(red, "kind", "") == "c10d":
            # torch.distributed collectives inside a captured step are not supported by this trainer (the NCCL
            # path hung under capture on B200); the NVLink backend's kernels are plain launches and capture fine
            self.cuda_graph = False
        if max_grad_norm is not None and hasattr(getattr(optimizer, "_symm", None), "_reduce_bucket"):
            raise ValueError("the bucket-fused sharded AdamW applies the update as each bucket's gradients arrive; "
                             "it cannot be combined with global gradient-norm clipping")
        # Optional (single-process CUDA training with this package's AdamW, no clipping): run the optimizer of the Linear
        # weights inside their wgrad GEMM epilogues — their gradients never reach HBM; everything else still goes
        # through step().  Off by default: on B200 the epilogue (26 B per accumulator element through 8 epilogue warps)
        # sustains 2.4 TB/s, the separate multi-tensor kernel 6 TB/s, so the fused GEMM is slower than GEMM + AdamW
        # (w1|w3 wgrad: 0.225 ms vs 0.160 ms, profiles/r2/gemm_vs_cublas_xl.json; XL step 48.7 vs 41.8 ms).
        self.optimizer_in_backward = False
        want = optimizer_in_backward
        if want is None:
            want = os.environ.get("B200SYS_OPT_IN_BWD", "0") == "1"
        if want and self.device.type == "cuda" and not self._is_ddp and max_grad_norm is None and \
                hasattr(optimizer, "fuse_into_backward") and autocast:
            self.optimizer_in_backward = optimizer.fuse_into_backward(model) > 0
        elif optimizer_in_backward:
            raise ValueError("optimizer_in_backward needs single-process CUDA training under autocast with "
                             "cs336_a2_systems_b200.optim.AdamW and no gradient clipping")
        if self.optimizer_in_backward:
            overlap_optimizer = False  # what is left for step() (norm gains, embeddings) is a few microseconds
        self.arena = None
        if grad_arena or (grad_arena is None and self.device.type == "cuda" and not self._is_ddp):
            # lay w3's slot next to w1's so their wgrads become ONE GEMM (B200SYS_FUSE_W13_WGRAD=0: parameter order)
            fuse_w13 = os.environ.get("B200SYS_FUSE_W13_WGRAD", "1") == "1"
            order = grad_slots.fusion_friendly_order(model) if fuse_w13 else list(model.parameters())
            # weights updated inside their wgrad GEMM never materialise a gradient: no slot (and no memory) for them
            order = [p for p in order if getattr(p, "_b200_bwd_opt", None) is None]
            self.arena = GradArena(order) if any(p.requires_grad for p in order) else None
        self._overlap = None
        if overlap_optimizer and self.device.type == "cuda" and max_grad_norm is None and \
                hasattr(optimizer, "step_subset"):
            self._overlap = _OptimizerOverlap(model, optimizer, self.device, overlap_unit_mb, self._is_ddp)
        self._graph = None
        self._static_x = self._static_y = self._static_loss = None
        self._warm_steps = 0
        #: eager steps before the capture (>= 1: optimizer plans, weight shadows and allocator pools must exist)
        self.graph_warmup_steps = max(1, int(graph_warmup_steps))
        #: kernels of the native library recorded into the step graph / number of replays so far (bench accounting)
        self.graph_kernel_count = 0
        self.graph_replays = 0
        self.steps_done = 0

    # ------------------------------------------------------------------ one eager step
    def _ctx(self):
        if self.autocast and self.device.type == "cuda":
            return torch.autocast("cuda", dtype=torch.bfloat16)
        return contextlib.nullcontext()

    def _step_eager(self, x: Tensor, y: Tensor) -> Tensor:
        if self._is_ddp and hasattr(self.model, "on_train_batch_start"):
            self.model.on_train_batch_start()
        if self.arena is not None:
            self.arena.release()
        self.optimizer.zero_grad(set_to_none=True)
        with nvtx.range("forward"), self._ctx():
            logits = self.model(x)
            loss = cross_entropy(logits, y)
        with nvtx.range("backward"):
            loss.backward()
        if self._is_ddp:
            with nvtx.range("grad_sync"):
                self.model.finish_gradient_synchronization()
        if self._overlap is not None:
            self._overlap.join()
        if self.max_grad_norm is not None:
            clip_gradient(self.model.parameters(), self.max_grad_norm)
        with nvtx.range("optimizer"):
            self.optimizer.step()
        return loss.detach()

    # ------------------------------------------------------------------ CUDA-graph step
    def _capture(self, x: Tensor, y: Tensor) -> None:
        self._static_x = x.clone()
        self._static_y = y.clone()
        self._graph = torch.cuda.CUDAGraph()
        # private side stream for capture, as torch requires
        s = torch.cuda.Stream(device=self.device)
        # push pending hyper-parameter changes (LR schedule) to the device NOW: inside the capture the optimizers
        # refuse to copy, because a recorded copy would be replayed with stale staging contents on every step
        refresh = getattr(self.optimizer, "refresh_device_hyper", None)
        if refresh is not None:
            refresh()
        s.wait_stream(torch.cuda.current_stream(self.device))
        n0 = kernels_mod.native_launches()
        with torch.cuda.stream(s):
            self.optimizer.zero_grad(set_to_none=True)
            with torch.cuda.graph(self._graph, stream=s):
                self._static_loss = self._step_eager(self._static_x, self._static_y)
        self.graph_kernel_count = kernels_mod.native_launches() - n0
        torch.cuda.current_stream(self.device).wait_stream(s)

    def step(self, x: Tensor, y: Tensor) -> Tensor:
        """One optimisation step on device-resident ``x, y`` (int64 ``[B, S]``). Returns the loss tensor."""
        self.steps_done += 1
        if not self.cuda_graph or self.device.type != "cuda":
            return self._step_eager(x, y)
        if self._graph is None:
            if self._warm_steps < self.graph_warmup_steps:  # plans, shadows and allocator pools must exist before capture
                self._warm_steps += 1
                return self._step_eager(x, y)
            self._capture(x, y)
            # the capture itself does not execute; fall through to a replay for this step
        self._static_x.copy_(x, non_blocking=True)
        self._static_y.copy_(y, non_blocking=True)
        refresh = getattr(self.optimizer, "refresh_device_hyper", None)
        if refresh is not None:  # LR schedules: the replay reads lr / betas / ... from device memory
            refresh()
        self._graph.replay()
        self.graph_replays += 1
        kernels_mod.bump_param_epoch()  # the replayed optimizer kernels rewrote the parameters
        return self._static_loss

    def step_from_host(self, x_host: Tensor, y_host: Tensor) -> float:
        """End-to-end step: pinned-host batch → device, train step, loss back to the host."""
        x = x_host.to(self.device, non_blocking=True)
        y = y_host.to(self.device, non_blocking=True)
        return float(self.step(x, y).item())
