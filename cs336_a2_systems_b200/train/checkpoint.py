"""Checkpoint / resume.

Model format is the course library's (``cs336_basics/model.py:312-327``): a directory with
``model_config.json`` (constructor kwargs) and ``model.pt`` (plain ``state_dict``, ``_orig_mod.`` prefix
tolerated on load).  The reference stops there (no optimizer state, no step counter, no resume;
SURVEY.md §5.4); this module adds ``trainer_state.pt`` with the optimizer state (consolidated across
ranks when it is sharded), the step counter and RNG state, written by rank 0 only.
"""
from __future__ import annotations

import json
import os

import torch
import torch.distributed as dist

from ..parallel.process_group import world


def _unwrap(model):
    return model.module if hasattr(model, "module") and hasattr(model, "finish_gradient_synchronization") else model


def save_checkpoint(path: str, model, optimizer=None, step: int = 0, extra: dict | None = None) -> None:
    rank, ws = world()
    m = _unwrap(model)
    opt_state = None
    if optimizer is not None:
        if hasattr(optimizer, "sync_step_from_device"):
            optimizer.sync_step_from_device()
        if hasattr(optimizer, "consolidated_state_dict"):
            opt_state = {"sharded": optimizer.consolidated_state_dict(dst=0)}  # collective: every rank calls
        else:
            opt_state = {"plain": optimizer.state_dict()}
    if rank == 0:
        os.makedirs(path, exist_ok=True)
        if hasattr(m, "config"):
            with open(os.path.join(path, "model_config.json"), "w") as f:
                json.dump(m.config, f)
        torch.save({k: v.detach().cpu() for k, v in m.state_dict().items()}, os.path.join(path, "model.pt"))
        torch.save({"step": step, "optimizer": opt_state, "extra": extra or {}, "rng": torch.get_rng_state()},
                   os.path.join(path, "trainer_state.pt"))
    if ws > 1:
        dist.barrier()


def load_checkpoint(path: str, model, optimizer=None, map_location="cpu") -> int:
    """Restores model (+ optimizer) in place on every rank; returns the saved step."""
    m = _unwrap(model)
    sd = torch.load(os.path.join(path, "model.pt"), map_location=map_location)
    prefix = "_orig_mod."
    sd = {(k[len(prefix):] if k.startswith(prefix) else k): v for k, v in sd.items()}
    with torch.no_grad():
        for k, v in m.state_dict().items():
            v.copy_(sd[k])  # in place: keeps arena-backed / symmetric parameter storage
    ts_path = os.path.join(path, "trainer_state.pt")
    if not os.path.exists(ts_path):
        return 0
    ts = torch.load(ts_path, map_location=map_location, weights_only=False)
    if optimizer is not None and ts.get("optimizer"):
        o = ts["optimizer"]
        if "sharded" in o and hasattr(optimizer, "load_consolidated_state_dict"):
            optimizer.load_consolidated_state_dict(o["sharded"])
        elif "plain" in o:
            optimizer.load_state_dict(o["plain"])
    return int(ts.get("step", 0))
