"""End-to-end LM on the native sm_100a path vs. the PyTorch oracle path (same weights)."""
import copy
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

from cs336_a2_systems_b200.models import TransformerLM  # noqa: E402
from cs336_a2_systems_b200.ops import cross_entropy  # noqa: E402
from cs336_a2_systems_b200.optim import AdamW  # noqa: E402


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


def _model(d_model=256, heads=4, layers=2, d_ff=512, vocab=1000, ctx=256):
    torch.manual_seed(0)
    return TransformerLM(vocab_size=vocab, context_length=ctx, d_model=d_model, num_layers=layers,
                         num_heads=heads, d_ff=d_ff, rope_theta=10000.0).cuda()


def _oracle_step(model, x, y):
    os.environ["B200SYS_FORCE_TORCH"] = "1"
    try:
        logits = model(x)  # fp32, plain PyTorch ops
        loss = torch.nn.functional.cross_entropy(logits.view(-1, logits.shape[-1]).float(), y.view(-1))
        loss.backward()
    finally:
        os.environ["B200SYS_FORCE_TORCH"] = "0"
    return logits.detach(), loss.detach()


@pytest.mark.parametrize("cfg", [dict(), dict(d_model=320, heads=4, d_ff=832)])  # head dim 64 and 80
def test_native_lm_matches_oracle(cfg):
    native = _model(**cfg)
    oracle = copy.deepcopy(native)
    x = torch.randint(0, 1000, (2, 256), device="cuda")
    y = torch.randint(0, 1000, (2, 256), device="cuda")
    ref_logits, ref_loss = _oracle_step(oracle, x, y)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        logits = native(x)
        loss = cross_entropy(logits, y)
    loss.backward()
    assert logits.dtype == torch.bfloat16
    assert _rel(logits, ref_logits) < 3e-2
    assert abs(float(loss) - float(ref_loss)) < 3e-2
    for (n, p), q in zip(native.named_parameters(), oracle.parameters()):
        assert p.grad is not None, n
        assert p.grad.dtype == torch.float32
        assert _rel(p.grad, q.grad) < 8e-2, f"{n}: {_rel(p.grad, q.grad)}"


def test_native_training_reduces_loss_and_refreshes_shadows():
    model = _model(layers=2)
    opt = AdamW(model.parameters(), lr=3e-3, weight_decay=0.0)
    x = torch.randint(0, 1000, (4, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)
    losses = []
    for _ in range(12):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = cross_entropy(model(x), y)
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0] - 0.5, losses
    # the fused optimizer keeps the bf16 GEMM operands in sync with the fp32 master weights
    w = model.layers[0].ffn.w2.weight
    from cs336_a2_systems_b200.ops import native
    (sh,) = native.shadow_slices_of(w)
    assert torch.equal(sh, w.detach().to(torch.bfloat16))


def test_gradients_land_in_registered_slots():
    from cs336_a2_systems_b200.ops import grad_slots
    model = _model(layers=1)
    params = [p for p in model.parameters()]
    total = sum((p.numel() + 3) // 4 * 4 for p in params)
    arena = torch.zeros(total, device="cuda")
    off = 0
    for p in params:
        grad_slots.register(p, arena, off)
        off += (p.numel() + 3) // 4 * 4
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    with torch.autocast("cuda", dtype=torch.bfloat16):
        cross_entropy(model(x), x).backward()
    lo, hi = arena.data_ptr(), arena.data_ptr() + arena.numel() * 4
    for n, p in model.named_parameters():
        assert lo <= p.grad.data_ptr() < hi, f"{n} gradient was not produced into its slot"
        grad_slots.unregister(p)


@pytest.mark.parametrize("graph", [False, True])
def test_trainer_overlapped_optimizer_matches_plain(graph):
    """Optimizer-in-backward (side stream) and the CUDA-graph step give the same trajectory as the plain loop."""
    from cs336_a2_systems_b200.train import Trainer
    base = _model(layers=2)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ta = Trainer(a, AdamW(a.parameters(), lr=1e-3, weight_decay=0.1), cuda_graph=False, overlap_optimizer=False)
    tb = Trainer(b, AdamW(b.parameters(), lr=1e-3, weight_decay=0.1), cuda_graph=graph, overlap_optimizer=True,
                 overlap_unit_mb=0.5)
    g = torch.Generator().manual_seed(3)
    for step in range(8):
        x = torch.randint(0, 1000, (2, 128), generator=g).cuda()
        y = torch.roll(x, -1, 1)
        la, lb = ta.step(x, y), tb.step(x, y)
        assert abs(float(la) - float(lb)) < 1e-5, (step, float(la), float(lb))
    torch.cuda.synchronize()
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        torch.testing.assert_close(p, q, rtol=1e-6, atol=1e-7, msg=n)
    tb.optimizer.sync_step_from_device()
    assert tb.optimizer.state[next(b.parameters())]["t"] == 9


def test_no_grad_forward_uses_gated_gemm_and_tracks_updates():
    """Under ``torch.no_grad()`` the FFN runs w1‖w3 as one GEMM with the SwiGLU gate in its epilogue (interleaved
    weight shadow).  It must agree with the training-mode forward, also after fused optimizer steps that rewrite
    the parameters without touching ``Tensor._version``."""
    model = _model(layers=2)
    opt = AdamW(model.parameters(), lr=3e-3, weight_decay=0.0)
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)

    def both():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            a = model(x).float()
            with torch.no_grad():
                b = model(x).float()
        return a.detach(), b

    a, b = both()
    assert _rel(b, a) < 2e-2
    assert hasattr(model.layers[0].ffn.w1.weight, "_b200_gate_shadow")
    before = b.clone()
    for _ in range(3):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            loss = cross_entropy(model(x), y)
        loss.backward()
        opt.step()
    a, b = both()
    assert _rel(b, a) < 2e-2
    assert _rel(b, before) > 1e-3  # the eval path saw the updated weights


def test_lr_schedule_reaches_the_captured_step():
    """Hyper-parameters live in device memory; the trainer refreshes them before the capture and before every
    CUDA-graph replay, so an LR schedule keeps working under ``cuda_graph=True``.  The learning rate changes on EVERY
    step here — the capture step included — so a staging copy recorded into the graph (which a replay would re-run
    with stale contents) shows up: with lr = 0 a replayed step must leave the parameters alone."""
    from cs336_a2_systems_b200.train import Trainer
    model = _model(layers=2)
    opt = AdamW(model.parameters(), lr=1e-3, weight_decay=0.1)
    tr = Trainer(model, opt, autocast=True, cuda_graph=True)
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)
    for i in range(6):  # eager warm-up steps, capture, replays
        opt.param_groups[0]["lr"] = 1e-3 * (1.0 + 0.1 * i)
        tr.step(x, y)
    assert tr._graph is not None
    for rounds in range(2):  # twice: the second round has wrapped further round the pinned staging ring
        torch.cuda.synchronize()
        before = [p.detach().clone() for p in model.parameters()]
        opt.param_groups[0]["lr"] = 0.0
        for _ in range(3):
            tr.step(x, y)
        torch.cuda.synchronize()
        for a, p in zip(before, model.parameters()):
            assert torch.equal(a, p.detach())
        opt.param_groups[0]["lr"] = 1e-3
        tr.step(x, y)
        torch.cuda.synchronize()
        assert any(not torch.equal(a, p.detach()) for a, p in zip(before, model.parameters()))


def _train(model, opt, x, y, steps):
    out = None
    for _ in range(steps):
        opt.zero_grad()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
            loss = cross_entropy(out, y)
        loss.backward()
        opt.step()
    return out.detach().float()


def test_unfused_adamw_refreshes_the_gemm_shadows():
    """The per-tensor PyTorch path of AdamW must invalidate the bf16 weight shadows the native GEMMs read: three
    unfused steps have to change the logits and land where three fused steps land."""
    a = _model(layers=2)
    b = copy.deepcopy(a)
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        first = a(x).float()
    la = _train(a, AdamW(a.parameters(), lr=3e-3, weight_decay=0.01, fused=False), x, y, 3)
    lb = _train(b, AdamW(b.parameters(), lr=3e-3, weight_decay=0.01, fused=True), x, y, 3)
    assert _rel(la, first) > 1e-2, "unfused AdamW steps did not reach the GEMM weights"
    assert _rel(la, lb) < 2e-2
    for p, q in zip(a.parameters(), b.parameters()):
        assert _rel(p, q) < 1e-3


def test_adamw_load_state_dict_into_a_live_optimizer():
    """step, save, step, load, step must equal a fresh optimizer resumed from the same checkpoint (the fused plan
    must not keep pointing at the moments that ``load_state_dict`` replaced, nor at its old step counter)."""
    a = _model(layers=1)
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)
    opt = AdamW(a.parameters(), lr=3e-3, weight_decay=0.01)
    _train(a, opt, x, y, 2)
    ck_model = copy.deepcopy(a.state_dict())
    ck_opt = copy.deepcopy(opt.state_dict())
    _train(a, opt, x, y, 2)          # moves on ...
    a.load_state_dict(ck_model)      # ... and rolls back
    opt.load_state_dict(copy.deepcopy(ck_opt))  # load_state_dict adopts the tensors it is given: keep ours pristine
    _train(a, opt, x, y, 1)
    b = _model(layers=1)
    b.load_state_dict(ck_model)
    opt_b = AdamW(b.parameters(), lr=3e-3, weight_decay=0.01)
    opt_b.load_state_dict(copy.deepcopy(ck_opt))
    _train(b, opt_b, x, y, 1)
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        d = (p - q).abs()
        # a step taken with stale moments / a stale step counter is off by ~lr = 3e-3 almost everywhere
        assert float((d > 2e-5 + 1e-4 * q.abs()).float().mean()) < 1e-3, \
            f"{n}: max diff {float(d.max()):.3e}, {int((d > 2e-5).sum())}/{d.numel()} elements off"
    for p, q in zip(a.parameters(), b.parameters()):
        assert opt.state[p]["t"] == opt_b.state[q]["t"]
        assert _rel(opt.state[p]["m"], opt_b.state[q]["m"]) < 1e-3


def test_optimizer_in_backward_matches_step_after_backward():
    """AdamW applied inside the wgrad GEMM epilogues (``Trainer(optimizer_in_backward=True)``) must train exactly like
    AdamW applied by ``optimizer.step()`` after backward: same arithmetic on the same fp32 accumulators — only the
    run-to-run rounding of FlashAttention's dQ reduction separates the two models.  Eager and CUDA-graph trainers."""
    from cs336_a2_systems_b200.train import Trainer
    x = torch.randint(0, 1000, (2, 128), device="cuda")
    y = torch.roll(x, -1, dims=1)
    for graph in (False, True):
        a, b = _model(layers=2), _model(layers=2)
        oa = AdamW(a.parameters(), lr=2e-3, weight_decay=0.1, betas=(0.9, 0.95))
        ob = AdamW(b.parameters(), lr=2e-3, weight_decay=0.1, betas=(0.9, 0.95))
        ta = Trainer(a, oa, autocast=True, cuda_graph=graph, optimizer_in_backward=True)
        tb = Trainer(b, ob, autocast=True, cuda_graph=graph, optimizer_in_backward=False)
        assert ta.optimizer_in_backward and not tb.optimizer_in_backward
        la = [float(ta.step(x, y)) for _ in range(6)]
        lb = [float(tb.step(x, y)) for _ in range(6)]
        torch.cuda.synchronize()
        assert la[-1] < la[0] and abs(la[-1] - lb[-1]) < 2e-2, (la, lb)
        lin = {id(m.weight) for m in a.modules() if m.__class__.__name__ == "Linear"}
        for (n, p), q in zip(a.named_parameters(), b.parameters()):
            if id(p) in lin:
                assert p.grad is None, f"{n}: fused weights never materialise a gradient"
            d = (p - q).abs()
            # Adam turns a sign flip of a ~0 gradient (fp32 summation order of dQ) into a +-lr step: bound the worst
            # element by steps x lr and judge agreement by the fraction of outliers
            assert float(d.max()) <= 6 * 2e-3 + 1e-6, f"{n}: {float(d.max()):.3e}"
            assert float((d > 1e-4 + 2e-3 * q.abs()).float().mean()) < 1e-2, n
        oa.sync_step_from_device()
        assert oa.state[next(iter(a.parameters()))]["t"] == 7 and all(st["t"] == 7 for st in oa.state.values())
        # the optimizer state is an ordinary state dict: a fresh unfused optimizer resumes from it
        sd = oa.state_dict()
        oc = AdamW(a.parameters(), lr=2e-3, weight_decay=0.1, betas=(0.9, 0.95))
        oc.load_state_dict(sd)
