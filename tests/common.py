"""Shared pieces of the multi-process tests: deterministic data, two toy networks, rendezvous helpers and the
cross-rank equality check.

The protocol is the reference's (``tests/common.py:13-94``): two gloo processes train replicas that are compared
with one in-process model that sees the whole batch.  What the toy networks must exercise is spelled out in
``TOY_SPECS`` below; the regression data comes from a seeded generator rather than fixture files.
"""
from __future__ import annotations

import os
from collections import OrderedDict

import torch
import torch.distributed as dist
from torch import nn

_N_FEATURES, _N_TARGETS = 10, 5


def make_data(n: int = 20, d_in: int = _N_FEATURES, d_out: int = _N_TARGETS):
    """``n`` regression samples ``(x[n, d_in], y[n, d_out])``, identical in every process."""
    gen = torch.Generator().manual_seed(1234)
    x = torch.randn(n, d_in, generator=gen)
    y = torch.randn(n, d_out, generator=gen)
    return x, y


class _Chain(nn.Module):
    """ReLU MLP assembled from ``(name, d_in, d_out, bias)`` rows; the last layer has no activation."""

    def __init__(self, rows):
        super().__init__()
        self._order = []
        for name, d_in, d_out, bias in rows:
            self.add_module(name, nn.Linear(d_in, d_out, bias=bias))
            self._order.append(name)

    def forward(self, x):
        last = len(self._order) - 1
        for i, name in enumerate(self._order):
            x = getattr(self, name)(x)
            if i != last:
                x = torch.relu(x)
        return x


class ToyModel(_Chain):
    """10 → 10 → 50 → 5.  The middle layer carries a bias that is frozen, and the model owns a frozen top-level
    parameter: DDP wrappers must skip ``requires_grad=False`` tensors when hooking and bucket accounting, yet still
    broadcast them."""

    def __init__(self):
        super().__init__([("fc1", _N_FEATURES, 10, False), ("fc2", 10, 50, True), ("fc3", 50, _N_TARGETS, False)])
        self.fc2.bias.requires_grad_(False)
        self.no_grad_fixed_param = nn.Parameter(torch.full((2,), 2.0), requires_grad=False)


class ToyModelWithTiedWeights(_Chain):
    """10 → 10 → 50 → 10 → 50 → 5 where the 4th layer re-uses the 2nd layer's weight: one parameter, two uses — a
    gradient hook must fire once for it, after both contributions have been accumulated."""

    def __init__(self):
        super().__init__([("fc1", _N_FEATURES, 10, False), ("fc2", 10, 50, False), ("fc3", 50, 10, False),
                          ("fc4", 10, 50, False), ("fc5", 50, _N_TARGETS, False)])
        self.fc4.weight = self.fc2.weight


TOY_SPECS = OrderedDict(
    frozen=("ToyModel", "frozen bias inside a layer + frozen top-level parameter"),
    tied=("ToyModelWithTiedWeights", "one weight shared by two layers"),
)


def assert_replicas_equal(net: nn.Module) -> None:
    """Collective: every entry of the wrapped module's state dict is the same on all ranks."""
    world = dist.get_world_size()
    for key, mine in net.module.state_dict().items():
        copies = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(copies, mine)
        for r, theirs in enumerate(copies):
            assert torch.allclose(theirs, mine), f"{key} differs between this rank and rank {r}"


def setup_process_group(rank: int, world_size: int, backend: str, port: int) -> str:
    """Rendezvous on 127.0.0.1:``port`` and return the device string this rank should use.

    As in the reference's helper a CUDA box maps rank r to ``cuda:(r mod #gpus)`` whatever the backend — with one GPU
    both gloo ranks land on ``cuda:0``, which is exactly the case where the wrappers have to fall back from the NVLink
    kernels to plain ``torch.distributed`` collectives."""
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    device = "cpu"
    n_gpu = torch.cuda.device_count() if torch.cuda.is_available() else 0
    if n_gpu:
        index = rank % n_gpu
        torch.cuda.set_device(index)
        device = f"cuda:{index}"
    dist.init_process_group(backend, rank=rank, world_size=world_size)
    return device


def cleanup_process_group() -> None:
    dist.barrier()
    dist.destroy_process_group()
