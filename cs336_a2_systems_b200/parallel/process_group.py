"""Process-group plumbing: one process per GPU, ``torch.distributed`` for rendezvous.

Counterpart of the reference's ad-hoc ``setup`` helpers (``ddp_bucketed_overlapped_sharded.py:46-49``,
``naive_ddp.py:35-51``, ``distributed_communication_single.py:10-26``): fixed ports and a
hard-coded world size of 2 there; env-driven (torchrun) or explicit here, always 127.0.0.1.
"""

from __future__ import annotations

import os
import socket

import torch
import torch.distributed as dist


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def setup(rank: int | None = None, world_size: int | None = None, backend: str | None = None,
          port: int | None = None, device: str | None = None) -> torch.device:
    """Initialise the default process group and return this rank's device.

    With no arguments the torchrun environment (``RANK``/``WORLD_SIZE``/``LOCAL_RANK``/``MASTER_*``)
    is used.  ``backend`` defaults to ``nccl`` when every rank can have its own GPU, else ``gloo``.
    """
    rank = int(os.environ.get("RANK", 0)) if rank is None else rank
    world_size = int(os.environ.get("WORLD_SIZE", 1)) if world_size is None else world_size
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if port is not None:
        os.environ["MASTER_PORT"] = str(port)
    os.environ.setdefault("MASTER_PORT", "29531")
    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if device is None:
        if n_gpu > 0:
            local = int(os.environ.get("LOCAL_RANK", rank % n_gpu)) % n_gpu
            torch.cuda.set_device(local)
            dev = torch.device("cuda", local)
        else:
            dev = torch.device("cpu")
    else:
        dev = torch.device(device)
        if dev.type == "cuda":
            torch.cuda.set_device(dev)
    if backend is None:
        backend = "nccl" if (dev.type == "cuda" and n_gpu >= world_size) else "gloo"
    if not dist.is_initialized():
        kw = {}
        if backend == "nccl":
            kw["device_id"] = dev
        dist.init_process_group(backend, rank=rank, world_size=world_size, **kw)
    return dev


def cleanup() -> None:
    if dist.is_initialized():
        try:
            dist.barrier()
        finally:
            dist.destroy_process_group()


def world() -> tuple[int, int]:
    """``(rank, world_size)``; ``(0, 1)`` when no process group exists (single-process use)."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1
