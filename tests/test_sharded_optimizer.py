"""Sharded optimizer == plain optimizer to default numpy tolerance (reference tests/test_sharded_optimizer.py)."""
import pytest
import torch.multiprocessing as mp

from cs336_a2_systems_b200.parallel import find_free_port

from ._ddp_worker import sharded_worker
from .common import ToyModel, ToyModelWithTiedWeights


@pytest.mark.parametrize("model_cls", [ToyModel, ToyModelWithTiedWeights])
def test_sharded_optimizer(model_cls):
    world_size = 2
    mp.spawn(sharded_worker, args=(world_size, find_free_port(), model_cls), nprocs=world_size, join=True)
