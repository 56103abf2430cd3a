"""Shared helpers for the distributed tests: toy models, process-group setup, equivalence checks.

Same protocol as the reference's ``tests/common.py:13-94`` (2-process gloo runs checked against a
single-process model trained on the full batch); data is generated from a fixed seed instead of
being loaded from fixture files.
"""
from __future__ import annotations

import os

import torch
import torch.distributed as dist
import torch.nn as nn


def make_data(n: int = 20, d_in: int = 10, d_out: int = 5):
    g = torch.Generator().manual_seed(1234)
    return torch.randn(n, d_in, generator=g), torch.randn(n, d_out, generator=g)


def assert_replicas_equal(net: nn.Module):
    """Every tensor of the wrapped module's state dict must be identical on all ranks."""
    for t in net.module.state_dict().values():
        gathered = [torch.zeros_like(t) for _ in range(dist.get_world_size())]
        dist.all_gather(gathered, t)
        for other in gathered:
            assert torch.allclose(other, t)


class _FrozenBiasFC(nn.Module):
    def __init__(self):
        super().__init__()
        self.fc = nn.Linear(10, 50, bias=True)
        self.fc.bias.requires_grad = False

    def forward(self, x):
        return self.fc(x)


class ToyModel(nn.Module):
    """Has a frozen bias and a frozen top-level parameter (requires_grad=False handling)."""

    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(10, 10, bias=False)
        self.fc2 = _FrozenBiasFC()
        self.fc3 = nn.Linear(50, 5, bias=False)
        self.relu = nn.ReLU()
        self.no_grad_fixed_param = nn.Parameter(torch.tensor([2.0, 2.0]), requires_grad=False)

    def forward(self, x):
        return self.fc3(self.relu(self.fc2(self.relu(self.fc1(x)))))


class ToyModelWithTiedWeights(nn.Module):
    """fc4 shares fc2's weight: the gradient hook must fire once per *parameter*."""

    def __init__(self):
        super().__init__()
        self.fc1 = nn.Linear(10, 10, bias=False)
        self.fc2 = nn.Linear(10, 50, bias=False)
        self.fc3 = nn.Linear(50, 10, bias=False)
        self.fc4 = nn.Linear(10, 50, bias=False)
        self.fc5 = nn.Linear(50, 5, bias=False)
        self.fc4.weight = self.fc2.weight
        self.relu = nn.ReLU()

    def forward(self, x):
        x = self.relu(self.fc1(x))
        x = self.relu(self.fc2(x))
        x = self.relu(self.fc3(x))
        x = self.relu(self.fc4(x))
        return self.fc5(x)


def setup_process_group(rank: int, world_size: int, backend: str, port: int) -> str:
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    if torch.cuda.is_available() and torch.cuda.device_count() > 0:
        # like the reference's helper: tensors go to cuda:(rank % n) even under gloo, so on a 1-GPU
        # box both ranks share cuda:0 and the wrappers must fall back to torch.distributed
        local = rank % torch.cuda.device_count()
        torch.cuda.set_device(local)
        device = f"cuda:{local}"
    else:
        device = "cpu"
    dist.init_process_group(backend, rank=rank, world_size=world_size)
    return device


def cleanup_process_group():
    dist.barrier()
    dist.destroy_process_group()
