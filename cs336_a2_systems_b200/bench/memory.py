"""Memory profiling: peak allocation and caching-allocator snapshots of a forward pass or a full step.

Counterpart of ``run_memory_foward_fullstep`` / ``profile_memory_forward_fullstep`` /
``profile_2_7b_memory_by_context_length`` (``cs336_systems/benchmark.py:175-245,314-447``): one forward or
one full train step under optional bf16 autocast with ``torch.cuda.memory._record_memory_history``,
``_dump_snapshot`` to ``memory_ctx{N}_{forward|fullstep}_{mixed|}.pickle`` (loadable by pytorch.org/memory_viz)
and ``max_memory_allocated``.  Default sweep = the reference's: 2.7B model × ctx {128,256,512}.

    python -m cs336_a2_systems_b200.bench.memory --size 2.7b --ctx 128 256 512 --out memory_files
"""
from __future__ import annotations

import argparse
import contextlib
import json
import os

import torch

from ..models import build_lm
from ..ops import cross_entropy
from ..optim import AdamW


def run_memory_forward_fullstep(size: str, ctx: int, full_step: bool, mixed: bool, batch: int = 4, vocab: int = 10000,
                                out_dir: str | None = None, snapshot: bool = True, device="cuda") -> dict:
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats()
    model = build_lm(size, vocab_size=vocab, context_length=ctx, device=device)
    opt = AdamW(model.parameters(), lr=1e-4) if full_step else None
    x = torch.randint(0, vocab, (batch, ctx), device=device)
    y = torch.roll(x, -1, 1)
    ac = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if mixed else contextlib.nullcontext
    if snapshot:
        torch.cuda.memory._record_memory_history(max_entries=1_000_000)
    if full_step:
        opt.zero_grad()
        with ac():
            loss = cross_entropy(model(x), y)
        loss.backward()
        opt.step()
    else:
        with torch.no_grad(), ac():
            model(x)
    torch.cuda.synchronize()
    peak = torch.cuda.max_memory_allocated() / 2 ** 20
    path = None
    if snapshot:
        if out_dir:
            os.makedirs(out_dir, exist_ok=True)
            path = os.path.join(out_dir, f"memory_ctx{ctx}_{'fullstep' if full_step else 'forward'}_{'mixed' if mixed else ''}.pickle")
            torch.cuda.memory._dump_snapshot(path)
        torch.cuda.memory._record_memory_history(enabled=None)
    del model, opt
    torch.cuda.empty_cache()
    return dict(size=size, ctx=ctx, phase="fullstep" if full_step else "forward", mixed=mixed, peak_mib=peak, snapshot=path)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--size", default="2.7b")
    ap.add_argument("--ctx", type=int, nargs="+", default=[128, 256, 512])
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--out", default=None, help="directory for snapshot pickles (omit: peaks only)")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    rows = []
    for ctx in a.ctx:
        for full in (False, True):
            for mixed in (False, True):
                r = run_memory_forward_fullstep(a.size, ctx, full, mixed, a.batch, out_dir=a.out, snapshot=a.out is not None)
                rows.append(r)
                print(json.dumps(r), flush=True)
    if a.json:
        json.dump(rows, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
