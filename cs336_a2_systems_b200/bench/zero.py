"""Micro-benchmark of the bucket-fused ZeRO-1 kernel (``csrc/comm_adamw.cu: zero_adamw_kernel``) against the
two-step path it replaces: NVLink all-reduce of the bucket followed by a replicated AdamW over the bucket.

What the reference does for the same work: ``dist.all_reduce`` per bucket, divide, ``optimizer.step()`` and one
``dist.broadcast`` per parameter (``ddp_bucketed_overlapped_sharded.py:289-318,335-340``).

    torchrun --nproc-per-node N -m cs336_a2_systems_b200.bench.zero [--sizes-mb 100 1000]

Roofline printed per variant: bytes over NVLink per rank (max of in/out direction) at the measured unidirectional
NVLink rate of the all-reduce kernel on the same box, and HBM bytes at the measured copy bandwidth.
"""
from __future__ import annotations

import argparse
import ctypes
import json

import torch
import torch.distributed as dist

from .. import _native
from ..ops import kernels
from ..parallel import setup, symm
from ..utils.clocks import ClockSampler
from .allreduce import _time


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--sizes-mb", type=float, nargs="+", default=[100, 1000])
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--blocks", type=int, nargs="+", default=[0])
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    dev = setup()
    rank, ws = dist.get_rank(), dist.get_world_size()
    nb = int(max(args.sizes_mb) * 1024 * 1024)
    garena = symm.SymmArena(nb + 256, dev, None)
    parena = symm.SymmArena(nb + 256, dev, None)
    sarena = symm.SymmArena(nb // 2 + 256, dev, None)
    hyper = torch.tensor([1e-4, 0.9, 0.95, 1e-8, 0.1, 1.0, 0, 0], dtype=torch.float32, device=dev)
    step = torch.ones(1, dtype=torch.int32, device=dev)
    mcs = [garena.multicast_ptr, parena.multicast_ptr, sarena.multicast_ptr]
    mc = kernels.host_ptr_array(mcs) if all(mcs) else None
    lib = _native.lib()
    stream = torch.cuda.current_stream(dev)
    rows = []
    for mb in args.sizes_mb:
        n = int(mb * 1024 * 1024) // 4
        g = garena.view(0, n, torch.float32)
        p = parena.view(0, n, torch.float32)
        g.normal_()
        p.normal_()
        nvec = (n * 4 + 31) // 32
        per = (nvec + ws - 1) // ws
        m = torch.zeros(per * 8, dtype=torch.float32, device=dev)
        v = torch.zeros(per * 8, dtype=torch.float32, device=dev)
        res = {"size_mb": mb, "world_size": ws}

        sampler = ClockSampler(dev.index or 0, period_ms=50).__enter__()
        none8 = (ctypes.c_longlong * 8)()
        table = torch.tensor([[0, nvec]], dtype=torch.int64, device=dev)

        def fused(variant, shadow, blocks, defer=0):
            _native.check(lib.b200_zero_adamw_bucket(
                garena.peer_ptrs_c, parena.peer_ptrs_c, sarena.peer_ptrs_c if shadow else None, garena.signal_ptrs_c,
                mc, ctypes.c_longlong(0), ctypes.c_longlong(n), _native.p(m), _native.p(v),
                _native.p(hyper), _native.p(step), ctypes.c_int(0), ctypes.c_int(rank), ctypes.c_int(ws),
                ctypes.c_int(blocks), ctypes.c_int(0), ctypes.c_int(variant), ctypes.c_int(defer), ctypes.c_int(0),
                none8, none8, ctypes.c_void_p(stream.cuda_stream)), "zero_adamw_bucket")

        def gather(use_mc):
            _native.check(lib.b200_zero_param_gather(
                parena.peer_ptrs_c, garena.signal_ptrs_c, ctypes.c_ulonglong(parena.multicast_ptr if use_mc else 0),
                _native.p(table), ctypes.c_int(1), ctypes.c_int(rank), ctypes.c_int(ws), ctypes.c_int(0),
                ctypes.c_int(0), ctypes.c_void_p(stream.cuda_stream)), "zero_param_gather")

        names = {0: "p2p", 1: "nvlsld_p2pst", 2: "p2pld_nvlsst", 3: "nvls"}
        for blocks in args.blocks:
            for variant in ([0, 1, 2, 3] if mc is not None else [0]):
                for shadow in (True, False):
                    key = f"fused_{names[variant]}_{'shadow' if shadow else 'noshadow'}_b{blocks}"
                    res[key + "_ms"] = _time(lambda: fused(variant, shadow, blocks), args.warmup, args.iters, dev)
                # deferred fp32 gather: the bucket kernel moves the shadows only; the gather kernel is timed apart
                res[f"fused_{names[variant]}_defer_b{blocks}_ms"] = _time(lambda: fused(variant, True, blocks, 1),
                                                                         args.warmup, args.iters, dev)
        res["param_gather_p2p_ms"] = _time(lambda: gather(False), args.warmup, args.iters, dev)
        if mc is not None:
            res["param_gather_nvls_ms"] = _time(lambda: gather(True), args.warmup, args.iters, dev)
        sh_all = sarena.view(0, n, torch.bfloat16)
        res["local_shadow_cast_ms"] = _time(lambda: kernels.cast_bf16(p, sh_all), args.warmup, args.iters, dev)
        # the two-step path: all-reduce (auto algorithm) + replicated AdamW over the whole bucket
        res["allreduce_ms"] = _time(lambda: kernels.allreduce_avg(garena, 0, n, torch.float32, 65536, 0), args.warmup,
                                    args.iters, dev)
        fm, fv = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
        sh = sarena.view(0, n, torch.bfloat16)
        launch = kernels.AdamWLaunch([p], [g], [fm], [fv], [sh], step0=1, mode=0)
        launch.set_hyper(1e-4, 0.9, 0.95, 1e-8, 0.1)
        res["adamw_replicated_ms"] = _time(lambda: launch.launch(), args.warmup, args.iters, dev)
        res["two_step_ms"] = res["allreduce_ms"] + res["adamw_replicated_ms"]
        # bytes over NVLink per rank (receive direction dominates): reduced slice + everybody else's new params
        S = n * 4
        res["nvlink_in_gb_shadow"] = (S / ws * (1 if mc is not None and ws > 2 else (ws - 1)) + 1.5 * S * (ws - 1) / ws) / 1e9
        res["nvlink_in_gb_allreduce"] = (S / ws + S * (ws - 1) / ws) / 1e9
        sampler.__exit__()
        res["clocks"] = sampler.summary()
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
        del launch, fm, fv, m, v
    if rank == 0 and args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    kernels.comm_check()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
