"""End-to-end LM benchmark: forward, backward and full-training-step time per model size.

Counterpart of the reference's ``run_simple_benchmark`` / ``run_all_benchmarks`` / ``run_benchmark_suite``
(``cs336_systems/benchmark.py:27-173,262-312``): warm-up steps, then per step a timed forward, a timed
backward and a timed full step (zero_grad/forward/CE/backward/AdamW), mean and std, one row per model
size.  Differences: CUDA-event timing instead of ``timeit`` + synchronize, optional CUDA-graph step, and a
Markdown/JSON table instead of LaTeX.

    python -m cs336_a2_systems_b200.bench.lm --sizes small medium --ctx 256 --steps 10 --json out.json
"""
from __future__ import annotations

import argparse
import contextlib
import json
import statistics

import torch

from ..models import MODEL_CONFIGS, build_lm
from ..ops import cross_entropy
from ..optim import AdamW
from ..train import Trainer


def _time(fn, n):
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return statistics.fmean(ts), (statistics.pstdev(ts) if len(ts) > 1 else 0.0)


def run_simple_benchmark(size: str, batch: int = 4, ctx: int = 256, vocab: int = 10000, warmup: int = 2,
                         steps: int = 10, mixed: bool = True, cuda_graph: bool = False, device="cuda") -> dict:
    model = build_lm(size, vocab_size=vocab, context_length=ctx, device=device)
    opt = AdamW(model.parameters(), lr=1e-4, betas=(0.9, 0.95), weight_decay=0.1)
    trainer = Trainer(model, opt, autocast=mixed, cuda_graph=cuda_graph)
    x = torch.randint(0, vocab, (batch, ctx), device=device)
    y = torch.roll(x, -1, 1)
    ac = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if mixed else contextlib.nullcontext
    for _ in range(max(warmup, 1) + (4 if cuda_graph else 0)):
        trainer.step(x, y)

    def fwd():
        with torch.no_grad(), ac():
            model(x)

    state = {}

    def fwd_keep():
        with ac():
            state["loss"] = cross_entropy(model(x), y)

    def bwd():
        model.zero_grad(set_to_none=True)
        if trainer.arena is not None:
            trainer.arena.release()
        state["loss"].backward()

    f_mean, f_std = _time(fwd, steps)
    b_times = []
    for _ in range(steps):
        fwd_keep()
        m, _ = _time(bwd, 1)
        b_times.append(m)
    s_mean, s_std = _time(lambda: trainer.step(x, y), steps)
    res = dict(size=size, batch=batch, ctx=ctx, mixed=mixed, cuda_graph=cuda_graph,
               params_m=sum(p.numel() for p in model.parameters()) / 1e6,
               forward_ms=f_mean, forward_std=f_std, backward_ms=statistics.fmean(b_times),
               backward_std=statistics.pstdev(b_times) if len(b_times) > 1 else 0.0,
               full_step_ms=s_mean, full_step_std=s_std, tokens_per_s=batch * ctx / (s_mean / 1e3),
               peak_mem_gib=torch.cuda.max_memory_allocated() / 2 ** 30)
    del model, opt, trainer
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    return res


def markdown(rows) -> str:
    out = ["| size | params (M) | fwd ms | bwd ms | full step ms | tokens/s | peak GiB |", "|---|---|---|---|---|---|---|"]
    for r in rows:
        out.append(f"| {r['size']} | {r['params_m']:.1f} | {r['forward_ms']:.2f} ± {r['forward_std']:.2f} | "
                   f"{r['backward_ms']:.2f} ± {r['backward_std']:.2f} | {r['full_step_ms']:.2f} ± {r['full_step_std']:.2f} | "
                   f"{r['tokens_per_s']:.0f} | {r['peak_mem_gib']:.1f} |")
    return "\n".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes", nargs="+", default=list(MODEL_CONFIGS))
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--ctx", type=int, default=256)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--fp32", action="store_true", help="no autocast (PyTorch oracle path)")
    ap.add_argument("--graph", action="store_true")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    rows = []
    for s in a.sizes:
        r = run_simple_benchmark(s, a.batch, a.ctx, warmup=a.warmup, steps=a.steps, mixed=not a.fp32, cuda_graph=a.graph)
        rows.append(r)
        print(json.dumps(r), flush=True)
    print(markdown(rows))
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
