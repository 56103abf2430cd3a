"""Training drivers: single-process, naive / flat / overlapped-per-parameter / bucketed DDP, optional
optimizer-state sharding — with per-phase timing and peak-memory logging.

Counterparts (reference): ``train_single_process`` / ``train_process`` / ``train_process_ddp`` and the
``--distributed --ddp --ddp_bucketed --sharded`` CLI of ``ddp_bucketed_overlapped_sharded.py:20-214,366-419``;
``train_regular`` / ``train`` / ``train_ddp`` / ``train_ddp_flat`` / ``compare_results`` of
``naive_ddp.py:53-732``.  Differences: world size is whatever the launcher provides (the reference
hard-codes 2), data is synthetic (its pickles are not shipped), step times are device-timed and reduced
with MAX over ranks, and every mode shares one loop.

Launch multi-process runs with torchrun (one rank per GPU) or ``spawn(...)`` below (mp.spawn, as the
reference does).
"""
from __future__ import annotations

import time
from dataclasses import asdict, dataclass, field

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from ..models import TransformerLM, get_model_config
from ..ops import clip_gradient, cross_entropy
from ..optim import AdamW
from ..parallel import (DDPBucketed, DDPIndividualParameters, ShardedOptimizer, allreduce_gradients_flat,
                        allreduce_gradients_naive, broadcast_parameters, cleanup, find_free_port, setup, world)
from ..utils import nvtx
from .data import SyntheticBatches, set_seed


@dataclass
class HParams:
    """Defaults = the reference CLI's hard-coded small model (``…sharded.py:390-404``)."""
    model: str = "small"
    d_model: int | None = None
    d_ff: int | None = None
    num_layers: int | None = None
    num_heads: int | None = None
    vocab_size: int = 10000
    context_length: int = 128
    batch_size: int = 128          # GLOBAL batch, split across ranks
    theta: float = 10000.0
    lr: float = 1e-3
    betas: tuple = (0.9, 0.95)
    eps: float = 1e-8
    weight_decay: float = 0.1
    num_steps: int = 50
    max_grad_norm: float | None = None
    autocast: bool = True
    bucket_size_mb: float = 1000.0
    seed: int = 0

    def model_kwargs(self):
        cfg = get_model_config(self.model)
        for k in ("d_model", "d_ff", "num_layers", "num_heads"):
            if getattr(self, k) is not None:
                cfg[k] = getattr(self, k)
        return dict(vocab_size=self.vocab_size, context_length=self.context_length, rope_theta=self.theta, **cfg)


@dataclass
class RunStats:
    mode: str
    world_size: int
    step_ms: list = field(default_factory=list)
    comm_ms: list = field(default_factory=list)
    losses: list = field(default_factory=list)
    mem_after_init_mb: float = 0.0
    mem_before_step_mb: float = 0.0
    mem_after_step_mb: float = 0.0

    def summary(self) -> dict:
        s = asdict(self)
        n = max(len(self.step_ms) - 1, 1)
        s["mean_step_ms"] = sum(self.step_ms[1:]) / n if len(self.step_ms) > 1 else (self.step_ms or [0])[0]
        if self.comm_ms:
            s["mean_comm_ms"] = sum(self.comm_ms[1:]) / n if len(self.comm_ms) > 1 else self.comm_ms[0]
            s["comm_fraction"] = s["mean_comm_ms"] / max(s["mean_step_ms"], 1e-9)
        return s


def _peak_mb(dev) -> float:
    if dev.type != "cuda":
        return 0.0
    v = torch.tensor([torch.cuda.max_memory_allocated(dev) / 2 ** 20], device=dev)
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(v, op=dist.ReduceOp.SUM)
        v /= dist.get_world_size()
    return float(v.item())


class _Timer:
    """CUDA events on GPU, perf_counter on CPU; ``ms()`` is MAX over ranks."""

    def __init__(self, dev):
        self.dev = dev
        self.gpu = dev.type == "cuda"

    def start(self):
        if self.gpu:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()
        else:
            self.t0 = time.perf_counter()

    def stop(self) -> float:
        if self.gpu:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            e1.synchronize()
            ms = self.e0.elapsed_time(e1)
        else:
            ms = (time.perf_counter() - self.t0) * 1e3
        if dist.is_initialized() and dist.get_world_size() > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=self.dev if dist.get_backend() == "nccl" else "cpu")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms


MODES = ("single", "naive", "flat", "ddp", "ddp_bucketed")


def train(hp: HParams, mode: str = "single", sharded: bool = False, device=None, log_every: int = 10,
          verbose: bool = True) -> RunStats:
    """One training run in ``mode`` ∈ {single, naive, flat, ddp, ddp_bucketed}; call inside an initialised
    process group for the distributed modes."""
    assert mode in MODES, mode
    rank, ws = world()
    dev = torch.device(device) if device is not None else (
        torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
    set_seed(hp.seed + (0 if mode == "single" else rank))
    if dev.type == "cuda":
        torch.cuda.reset_peak_memory_stats(dev)
        with torch.device(dev):
            model = TransformerLM(**hp.model_kwargs())
    else:
        model = TransformerLM(**hp.model_kwargs())
    fwd = model
    if mode == "ddp":
        fwd = DDPIndividualParameters(model)
    elif mode == "ddp_bucketed":
        fwd = DDPBucketed(model, bucket_size_mb=hp.bucket_size_mb)
    elif mode in ("naive", "flat") and ws > 1:
        broadcast_parameters(model, 0)
    kw = dict(lr=hp.lr, betas=hp.betas, eps=hp.eps, weight_decay=hp.weight_decay)
    opt = ShardedOptimizer(model.parameters(), AdamW, **kw) if sharded else AdamW(model.parameters(), **kw)
    if sharded and dev.type == "cuda":
        opt.enable_symmetric()  # fused owner-AdamW + NVLink distribute when legal; no-op otherwise
    stats = RunStats(mode=mode + ("+sharded" if sharded else ""), world_size=ws)
    stats.mem_after_init_mb = _peak_mb(dev)
    data = SyntheticBatches(hp.batch_size, hp.context_length, hp.vocab_size, rank, ws, seed=hp.seed + 1)
    timer, comm_timer = _Timer(dev), _Timer(dev)
    ctx = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if (hp.autocast and dev.type == "cuda") else None
    for step in range(hp.num_steps):
        xh, yh = next(data)
        x, y = xh.to(dev, non_blocking=True), yh.to(dev, non_blocking=True)
        if dev.type == "cuda":
            torch.cuda.reset_peak_memory_stats(dev)
        timer.start()
        if hasattr(fwd, "on_train_batch_start"):
            fwd.on_train_batch_start()
        opt.zero_grad()
        if ctx is not None:
            with ctx():
                loss = cross_entropy(fwd(x), y)
        else:
            loss = cross_entropy(fwd(x), y)
        with nvtx.range("Backward"):
            loss.backward()
        if mode in ("naive", "flat") and ws > 1:
            comm_timer.start()
            with nvtx.range("Communication"):
                (allreduce_gradients_naive if mode == "naive" else allreduce_gradients_flat)(model)
            stats.comm_ms.append(comm_timer.stop())
        elif mode in ("ddp", "ddp_bucketed"):
            fwd.finish_gradient_synchronization()
        if hp.max_grad_norm is not None:
            clip_gradient(model.parameters(), hp.max_grad_norm)
        if step == hp.num_steps - 1:
            stats.mem_before_step_mb = _peak_mb(dev)
        opt.step()
        stats.step_ms.append(timer.stop())
        if step == hp.num_steps - 1:
            stats.mem_after_step_mb = _peak_mb(dev)
        if step % log_every == 0 or step == hp.num_steps - 1:
            l = loss.detach().clone()
            if ws > 1 and mode != "single":
                dist.all_reduce(l, op=dist.ReduceOp.SUM)
                l /= ws
            stats.losses.append(float(l))
            if verbose and rank == 0:
                print(f"[{stats.mode}] step {step} loss {float(l):.4f} step {stats.step_ms[-1]:.2f} ms", flush=True)
    if verbose and rank == 0:
        s = stats.summary()
        print(f"[{stats.mode}] mean step {s['mean_step_ms']:.2f} ms"
              + (f", comm {s['mean_comm_ms']:.2f} ms ({100 * s['comm_fraction']:.1f}%)" if stats.comm_ms else "")
              + f"; peak MB after init {stats.mem_after_init_mb:.0f}, before step {stats.mem_before_step_mb:.0f},"
                f" after step {stats.mem_after_step_mb:.0f}", flush=True)
    stats.final_state = {k: v.detach().cpu() for k, v in model.state_dict().items()}
    return stats


def compare_results(a: dict, b: dict, rtol: float = 1e-4, atol: float = 1e-5) -> dict:
    """Parameter-by-parameter comparison of two final state dicts (reference ``naive_ddp.py:636-658``)."""
    worst, bad = 0.0, []
    for k in a:
        d = float((a[k].float() - b[k].float()).abs().max())
        worst = max(worst, d)
        if not torch.allclose(a[k].float(), b[k].float(), rtol=rtol, atol=atol):
            bad.append(k)
    return {"max_abs_diff": worst, "mismatched": bad, "match": not bad}


def _worker(rank: int, world_size: int, port: int, hp: HParams, mode: str, sharded: bool, queue):
    dev = setup(rank, world_size, port=port)
    try:
        stats = train(hp, mode, sharded, device=dev, verbose=True)
        if rank == 0 and queue is not None:
            s = stats.summary()
            queue.put(s)
    finally:
        cleanup()


def spawn(hp: HParams, mode: str, sharded: bool = False, world_size: int = 2) -> dict:
    """mp.spawn launcher (what the reference's ``train_distributed*`` helpers do)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    mp.spawn(_worker, args=(world_size, find_free_port(), hp, mode, sharded, q), nprocs=world_size, join=True)
    return q.get()
