"""Bucketed overlapped DDP == single-process training (reference ``tests/test_ddp.py:29-181``).

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


@pytest.mark.parametrize("size,expected", [("small", [98, 50, 6, 1]), ("xl", [386, 339, 98, 8]),
                                           ("2.7b", [258, 258, 161, 14])])
def test_bucket_counts_match_reference_algorithm(size, expected):
    """The greedy reverse-order bucketing of the reference (``ddp_bucketed_overlapped_sharded.py:265-279``) yields
    these bucket counts for bucket_size_mb = 1/10/100/1000; ours must plan the same buckets."""
    import torch
    from cs336_a2_systems_b200.models import build_lm
    from cs336_a2_systems_b200.parallel.ddp import plan_buckets

    m = build_lm(size, vocab_size=10000, context_length=512, device="meta")
    order = list(reversed([p for p in m.parameters() if p.requires_grad]))
    assert [len(plan_buckets(order, mb)) for mb in (1, 10, 100, 1000)] == expected
    # every parameter lands in exactly one bucket
    for mb in (1, 100):
        idx = sorted(i for b in plan_buckets(order, mb) for i in b)
        assert idx == list(range(len(order)))
