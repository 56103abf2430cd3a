"""CLI: ``python -m cs336_a2_systems_b200.train [--distributed] [--ddp | --ddp_bucketed] [--sharded] …``

Same four switches as the reference CLI (``ddp_bucketed_overlapped_sharded.py:366-419``); everything the
reference hard-codes is a flag here.  Under torchrun (``WORLD_SIZE`` set) the existing ranks are used;
otherwise ``--distributed/--ddp/--ddp_bucketed`` spawn ``--world-size`` local processes.
"""
from __future__ import annotations

import argparse
import json
import os

from ..parallel import cleanup, setup
from .drivers import HParams, spawn, train


def main():
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--distributed", action="store_true", help="naive DDP: per-parameter all-reduce after backward")
    ap.add_argument("--flat", action="store_true", help="flat DDP: one all-reduce of all gradients")
    ap.add_argument("--ddp", action="store_true", help="overlapped per-parameter DDP")
    ap.add_argument("--ddp_bucketed", action="store_true", help="overlapped bucketed DDP")
    ap.add_argument("--sharded", action="store_true", help="optimizer-state sharding")
    ap.add_argument("--world-size", type=int, default=2)
    ap.add_argument("--model", default="small")
    ap.add_argument("--batch-size", type=int, default=128)
    ap.add_argument("--context-length", type=int, default=128)
    ap.add_argument("--num-steps", type=int, default=50)
    ap.add_argument("--bucket-size-mb", type=float, default=1000.0)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--max-grad-norm", type=float, default=None)
    ap.add_argument("--no-autocast", action="store_true")
    ap.add_argument("--json", default=None)
    a = ap.parse_args()
    hp = HParams(model=a.model, batch_size=a.batch_size, context_length=a.context_length, num_steps=a.num_steps,
                 bucket_size_mb=a.bucket_size_mb, lr=a.lr, max_grad_norm=a.max_grad_norm, autocast=not a.no_autocast)
    mode = "ddp_bucketed" if a.ddp_bucketed else "ddp" if a.ddp else "flat" if a.flat else "naive" if a.distributed else "single"
    if mode == "single":
        out = train(hp, "single", sharded=False).summary()
    elif int(os.environ.get("WORLD_SIZE", "1")) > 1:
        setup()
        try:
            out = train(hp, mode, a.sharded).summary()
        finally:
            cleanup()
    else:
        out = spawn(hp, mode, a.sharded, a.world_size)
    out.pop("final_state", None)
    if a.json and int(os.environ.get("RANK", "0")) == 0:
        json.dump(out, open(a.json, "w"), indent=1)


if __name__ == "__main__":
    main()
