"""Per-parameter overlapped DDP == single-process training (reference
``tests/test_ddp_individual_parameters.py:28-161``: 2-process gloo, rank-seeded replicas equalised by the wrapper, 5 SGD steps
on disjoint shards vs one model on the full batch)."""
import pytest
import torch.multiprocessing as mp

from cs336_a2_systems_b200.parallel import find_free_port

from ._ddp_worker import ddp_worker
from .common import ToyModel, ToyModelWithTiedWeights


@pytest.mark.parametrize("model_cls", [ToyModel, ToyModelWithTiedWeights])
def test_ddp_individual_parameters(model_cls):
    world_size = 2
    mp.spawn(ddp_worker, args=(world_size, find_free_port(), model_cls, "individual", None),
             nprocs=world_size, join=True)
