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


def test_bucket_slot_layout_invariants():
    """Single process (no process group): every trainable parameter gets exactly one slot, slots are 32-byte aligned,
    do not overlap, follow forward parameter order inside a bucket (so q/k/v gradients are adjacent and one wgrad
    GEMM can produce all three)."""
    import torch
    from cs336_a2_systems_b200.models import TransformerLM
    from cs336_a2_systems_b200.parallel import DDPBucketed

    torch.manual_seed(0)
    m = TransformerLM(vocab_size=50, context_length=16, d_model=32, num_layers=2, num_heads=2, d_ff=64)
    ddp = DDPBucketed(m, bucket_size_mb=0.02)
    assert len(ddp.buckets) > 2
    seen = set()
    order = {id(p): i for i, p in enumerate(m.parameters())}
    for b in ddp.buckets:
        idx = [order[id(p)] for p in b.params]
        assert idx == sorted(idx), "slots inside a bucket follow forward parameter order"
        end = 0
        for p, off in zip(b.params, b.offsets):
            assert id(p) not in seen
            seen.add(id(p))
            assert off >= end and (off * p.element_size()) % DDPBucketed.ALIGN_BYTES == 0
            end = off + p.numel()
        assert end <= b.numel == b.flat.numel()
    assert seen == {id(p) for p in m.parameters() if p.requires_grad}
    blk = m.layers[0].attn
    where = {id(p): (bi, off) for bi, b in enumerate(ddp.buckets) for p, off in zip(b.params, b.offsets)}
    (bq, oq), (bk, ok_), (bv, ov) = (where[id(w.weight)] for w in (blk.q_proj, blk.k_proj, blk.v_proj))
    if bq == bk == bv:
        assert ok_ == oq + blk.q_proj.weight.numel() and ov == ok_ + blk.k_proj.weight.numel()
    # with a single process there is nothing to reduce: the wrapper must still train like the bare module
    x = torch.randint(0, 50, (2, 16))
    ddp.on_train_batch_start()
    ddp(x).sum().backward()
    ddp.finish_gradient_synchronization()
    assert all(p.grad is not None for p in m.parameters())
