"""Worker bodies for the 2-process gloo DDP / sharded-optimizer tests (spawned with mp.spawn);
protocols of the reference's ``tests/test_ddp.py:29-181``, ``test_ddp_individual_parameters.py:39-161`` and
``test_sharded_optimizer.py:29-85``."""
from __future__ import annotations

import copy

import numpy
import torch
import torch.nn as nn

from . import adapters
from .common import (assert_replicas_equal, cleanup_process_group, make_data, setup_process_group)


def ddp_worker(rank: int, world_size: int, port: int, model_cls, mode: str, bucket_size_mb):
    device = setup_process_group(rank, world_size, "gloo", port)
    torch.manual_seed(rank)  # replicas start different on purpose
    base = model_cls().to(device)
    ref = copy.deepcopy(base)
    if mode == "individual":
        ddp = adapters.get_ddp_individual_parameters(base)
    else:
        ddp = adapters.get_ddp_bucketed(base, bucket_size_mb)
    # after wrapping, every rank holds rank 0's values
    for (n_ref, p_ref), (n_ddp, p_ddp) in zip(ref.named_parameters(), ddp.module.named_parameters()):
        assert n_ref == n_ddp
        same = torch.allclose(p_ref, p_ddp)
        is_const = "no_grad_fixed_param" in n_ddp
        if rank == 0 or is_const:
            assert same
        else:
            assert not same, f"{n_ddp} was not overwritten by rank 0's value"
    assert_replicas_equal(ddp)
    # wrapper adds no parameters of its own and keeps names aligned with the bare model
    assert [n.removeprefix("module.") for n, _ in ddp.named_parameters()] == [n for n, _ in base.named_parameters()]

    torch.manual_seed(42)
    single = model_cls().to(device)
    single.load_state_dict(ddp.module.state_dict())
    opt_ddp = torch.optim.SGD(ddp.parameters(), lr=0.1)
    opt_single = torch.optim.SGD(single.parameters(), lr=0.1)
    loss_fn = nn.MSELoss()
    x_all, y_all = (t.to(device) for t in make_data())
    shard = x_all.shape[0] // world_size
    for step in range(5):
        if mode == "bucketed":
            adapters.ddp_bucketed_on_train_batch_start(ddp, opt_ddp)
        opt_single.zero_grad()
        loss_fn(single(x_all), y_all).backward()
        opt_single.step()

        opt_ddp.zero_grad()
        xs = x_all[rank * shard:(rank + 1) * shard]
        ys = y_all[rank * shard:(rank + 1) * shard]
        loss_fn(ddp(xs), ys).backward()
        if mode == "individual":
            adapters.ddp_individual_parameters_on_after_backward(ddp, opt_ddp)
        else:
            adapters.ddp_bucketed_on_after_backward(ddp, opt_ddp)
        opt_ddp.step()

        for (n, p_s), p_d in zip(single.named_parameters(), ddp.module.parameters()):
            assert torch.allclose(p_s, p_d, atol=1e-6), f"step {step}: {n} diverged from single-process run"
            if not p_s.requires_grad:
                assert p_d.grad is None
        # reshuffle with a shared seed so ranks stay consistent
        g = torch.Generator().manual_seed(100 + step)
        perm = torch.randperm(x_all.shape[0], generator=g).to(device)
        x_all, y_all = x_all[perm], y_all[perm]
    assert_replicas_equal(ddp)
    cleanup_process_group()


def sharded_worker(rank: int, world_size: int, port: int, model_cls):
    device = setup_process_group(rank, world_size, "gloo", port)
    torch.manual_seed(42)
    plain = model_cls().to(device)
    sharded = copy.deepcopy(plain)
    kw = dict(lr=0.1, weight_decay=0.1, betas=(0.9, 0.999), eps=1e-8)
    opt_plain = torch.optim.AdamW(plain.parameters(), **kw)
    opt_sharded = adapters.get_sharded_optimizer(sharded.parameters(), torch.optim.AdamW, **kw)
    x, y = (t.to(device) for t in make_data())
    for _ in range(10):
        for model, opt in ((plain, opt_plain), (sharded, opt_sharded)):
            opt.zero_grad()
            ((model(x) - y) ** 2).sum().backward()
            opt.step()
    for p_a, p_b in zip(plain.parameters(), sharded.parameters()):
        numpy.testing.assert_allclose(p_a.detach().cpu().numpy(), p_b.detach().cpu().numpy())
    # state really is sharded: this rank holds state only for what it owns
    n_state = len(opt_sharded.optimizer.state) if opt_sharded.optimizer is not None else 0
    n_train = sum(1 for p in sharded.parameters() if p.requires_grad)
    assert 0 < n_state < n_train
    sd = opt_sharded.consolidated_state_dict(dst=0)
    if rank == 0:
        assert len(sd["state"]) == n_train
    cleanup_process_group()
