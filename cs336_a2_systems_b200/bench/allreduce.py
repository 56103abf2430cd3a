"""All-reduce micro-benchmark: our NVLink kernels (one-shot / two-shot / multimem) vs NCCL.

Counterpart of the reference's ``distributed_communication_single.py:28-151`` (fp32 tensors of
1 MB … 1 GB, warm-up then timed iterations, per-rank stats gathered), but device-timed with CUDA
events, max over ranks, and reporting algorithm + bus bandwidth.  Launch with torchrun:

    torchrun --nproc-per-node N -m cs336_a2_systems_b200.bench.allreduce [--sizes-mb 1 10 100 1000]
"""
from __future__ import annotations

import argparse
import json

import torch
import torch.distributed as dist

from ..ops import kernels
from ..parallel import setup, symm
from ..utils.clocks import ClockSampler
from ..utils.timing import time_collective


def _time(fn, warmup, iters, dev):
    return time_collective(fn, warmup, iters, dev)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", type=float, nargs="+", default=[0.008, 0.064, 0.5, 1, 10, 100, 1000])
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--blocks", type=int, default=0)
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev = setup()
    rank, ws = dist.get_rank(), dist.get_world_size()
    max_bytes = int(max(args.sizes_mb) * 1024 * 1024)
    arena = symm.SymmArena(max_bytes + 256, dev, None)
    algos = {"oneshot": 1, "twoshot": 2}
    if arena.multicast_ptr:
        algos["multimem"] = 3
    rows = []
    for mb in args.sizes_mb:
        n = int(mb * 1024 * 1024) // 4
        buf = arena.view(0, n, torch.float32)
        buf.normal_()
        ref = buf.clone()
        res = {"size_mb": mb, "world_size": ws}
        sampler = ClockSampler(dev.index or 0, period_ms=50).__enter__()
        t = _time(lambda: dist.all_reduce(ref), args.warmup, args.iters, dev)
        res["nccl_ms"] = t
        for name, algo in algos.items():
            if algo == 1 and n * 4 > 8 * 1024 * 1024:
                continue
            t = _time(lambda: kernels.allreduce_avg(arena, 0, n, torch.float32, one_shot_max=0, blocks=args.blocks,
                                                    algo=algo), args.warmup, args.iters, dev)
            res[f"{name}_ms"] = t
        # bus bandwidth (NCCL convention): 2 (ws-1)/ws * S / t
        for k in list(res):
            if k.endswith("_ms") and not k.startswith("bcast"):
                res[k.replace("_ms", "_busbw_gbs")] = 2 * (ws - 1) / ws * n * 4 / (res[k] * 1e-3) / 1e9
        if n * 4 >= 1 << 20:
            # parameter broadcast from rank 0: NCCL vs pull (ws-1 readers) vs one multicast store through the switch
            res["bcast_nccl_ms"] = _time(lambda: dist.broadcast(ref, src=0), args.warmup, args.iters, dev)
            res["bcast_pull_ms"] = _time(lambda: kernels.bcast_pull(arena, 0, n * 4, 0), args.warmup, args.iters, dev)
            if arena.multicast_ptr:
                res["bcast_mc_ms"] = _time(lambda: kernels.bcast_mc(arena, 0, n * 4, 0), args.warmup, args.iters, dev)
        sampler.__exit__()
        res["clocks"] = sampler.summary()  # rank 0's GPU while this row ran
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
    kernels.comm_check()
    if rank == 0 and args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
