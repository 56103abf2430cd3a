"""Bucketed overlapped DDP == single-process training (reference tests/test_ddp.py).

Bucket sizes are chosen so the toy models split into two buckets, one bucket per parameter, and a
single bucket.
"""
import pytest
import torch
import torch.multiprocessing as mp

from cs336_a2_systems_b200.parallel import find_free_port, plan_buckets

from ._ddp_worker import ddp_worker
from .common import ToyModel, ToyModelWithTiedWeights


@pytest.mark.parametrize("bucket_size_mb", [0.0016, 0.0001, 0.01])
@pytest.mark.parametrize("model_cls", [ToyModel, ToyModelWithTiedWeights])
def test_ddp_bucketed(model_cls, bucket_size_mb):
    world_size = 2
    mp.spawn(ddp_worker, args=(world_size, find_free_port(), model_cls, "bucketed", bucket_size_mb),
             nprocs=world_size, join=True)


def test_bucket_planner():
    ps = [torch.zeros(n) for n in (100, 100, 300, 50, 50)]          # 400, 400, 1200, 200, 200 bytes
    mb = 1.0 / (1024 * 1024)
    assert plan_buckets(ps, 800 * mb) == [[0, 1], [2], [3, 4]]       # oversize tensor gets its own bucket
    assert plan_buckets(ps, None) == [[0, 1, 2, 3, 4]]               # unbounded: one bucket
    assert plan_buckets(ps, 1 * mb) == [[0], [1], [2], [3], [4]]     # tiny cap: one per tensor
    mixed = [torch.zeros(4), torch.zeros(4, dtype=torch.float64), torch.zeros(4)]
    assert plan_buckets(mixed, None) == [[0], [1], [2]]              # dtype change closes a bucket
