"""Sharded optimizer == plain optimizer to default numpy tolerance (reference
``tests/test_sharded_optimizer.py:18-85``: torch.optim.AdamW plain vs sharded, 10 steps, ``assert_allclose`` at rtol 1e-7)."""
import pytest
import torch.multiprocessing as mp

from cs336_a2_systems_b200.parallel import find_free_port

from ._ddp_worker import sharded_worker
from .common import ToyModel, ToyModelWithTiedWeights


@pytest.mark.parametrize("model_cls", [ToyModel, ToyModelWithTiedWeights])
def test_sharded_optimizer(model_cls):
    world_size = 2
    mp.spawn(sharded_worker, args=(world_size, find_free_port(), model_cls), nprocs=world_size, join=True)


def test_zero_bucket_slice_arithmetic():
    """Pure bookkeeping of the bucket-fused ZeRO path (parallel/zero.py): rank slices tile the bucket, and
    per-parameter optimizer state survives assemble -> slice -> gather -> split."""
    import torch
    from cs336_a2_systems_b200.parallel.zero import assemble_bucket_state, slice_floats, split_bucket_state

    g = torch.Generator().manual_seed(0)
    shapes = [(5, 7), (3,), (16, 8), (1, 1)]
    offsets, numel = [], 0
    for sh in shapes:  # 32-byte aligned slots, like DDPBucketed
        offsets.append(numel)
        n = int(torch.tensor(sh).prod())
        numel += (n + 7) // 8 * 8
    for ws in (1, 2, 3, 4, 8):
        per8 = slice_floats(numel, ws)
        assert per8 % 8 == 0 and per8 * ws >= numel and per8 * (ws - 1) < numel + 8 * ws
        state = [torch.randn(sh, generator=g) for sh in shapes]
        full = assemble_bucket_state(state, offsets, per8 * ws)
        slices = [full[r * per8:(r + 1) * per8].clone() for r in range(ws)]  # what each rank keeps
        back = split_bucket_state(torch.cat(slices), shapes, offsets)       # what all_gather + split returns
        for a, b in zip(state, back):
            torch.testing.assert_close(a, b, rtol=0, atol=0)


def test_partition_is_greedy_least_loaded_and_total(monkeypatch):
    """Ownership follows the reference's rule (``ddp_bucketed_overlapped_sharded.py:342-347``: next parameter to the
    rank with the fewest bytes so far, ties to the lowest rank), every parameter has exactly one owner, every rank
    computes the same map, and the byte loads end up balanced for the LM's parameter mix."""
    import torch
    from cs336_a2_systems_b200.models import build_lm
    from cs336_a2_systems_b200.parallel import sharded_optimizer as so

    model = build_lm("small", vocab_size=1000, context_length=64, device="meta")
    params = list(model.parameters())
    ws = 4
    maps = []
    for rank in range(ws):
        monkeypatch.setattr(so, "world", lambda r=rank: (r, ws))
        opt = so.ShardedOptimizer(params, torch.optim.SGD, lr=0.1)
        maps.append([opt.owner_of(p) for p in params])
        local = [] if opt.optimizer is None else [p for g in opt.optimizer.param_groups for p in g["params"]]
        assert {id(p) for p in local} == {id(p) for p, o in zip(params, maps[-1]) if o == rank}
    assert all(m == maps[0] for m in maps)
    load = [0] * ws
    for p, o in zip(params, maps[0]):
        assert o == min(range(ws), key=lambda r: (load[r], r))
        load[o] += p.numel() * p.element_size()
    assert max(load) / min(load) < 1.2
