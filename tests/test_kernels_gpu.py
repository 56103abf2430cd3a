"""Numerics of every sm_100a kernel against a plain-PyTorch fp32 reference of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu

from cs336_a2_systems_b200.ops import functional as F_ref  # noqa: E402


def _k():
    from cs336_a2_systems_b200.ops import kernels
    return kernels


def _bf(t):
    return t.to(torch.bfloat16)


def _rel_err(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


# ------------------------------------------------------------------------------------------------ GEMM
@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (256, 256, 128), (2048, 1600, 1600), (2048, 4800, 1600),
                                   (2048, 10000, 1600), (200, 136, 72), (64, 48, 40), (1, 8, 8)])
@pytest.mark.parametrize("bn", [0, 128, 192, 256])
def test_gemm_nt(M, N, K, bn):
    k = _k()
    torch.manual_seed(0)
    a = _bf(torch.randn(M, K, device="cuda"))
    b = _bf(torch.randn(N, K, device="cuda"))
    c = k.gemm(a, b, force_bn=bn)
    ref = a.float() @ b.float().t()
    assert _rel_err(c, ref) < 6e-3
    c32 = k.gemm(a, b, epilogue=k.EPI_F32, force_bn=bn)
    torch.testing.assert_close(c32, ref, rtol=1e-3, atol=1e-2 * math.sqrt(K) / 8)


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (2048, 1600, 4800), (256, 192, 320), (200, 136, 72)])
def test_gemm_nn_dgrad(M, N, K):
    """dx = dy · W with W stored [K, N] (MN-major B)."""
    k = _k()
    torch.manual_seed(1)
    dy = _bf(torch.randn(M, K, device="cuda"))
    w = _bf(torch.randn(K, N, device="cuda"))
    out = k.gemm(dy, w, b_mn=True)
    assert _rel_err(out, dy.float() @ w.float()) < 6e-3


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (1600, 1600, 2048), (4800, 1600, 2048), (136, 200, 72)])
def test_gemm_tn_wgrad(M, N, K):
    """dW = dyᵀ · x with dy [K, M] and x [K, N] (both MN-major), fp32 output, plus accumulate epilogue."""
    k = _k()
    torch.manual_seed(2)
    dy = _bf(torch.randn(K, M, device="cuda"))
    x = _bf(torch.randn(K, N, device="cuda"))
    ref = dy.float().t() @ x.float()
    out = k.gemm(dy, x, a_mn=True, b_mn=True, epilogue=k.EPI_F32)
    assert _rel_err(out, ref) < 2e-3
    k.gemm(dy, x, a_mn=True, b_mn=True, epilogue=k.EPI_F32_ACC, out=out)
    assert _rel_err(out, 2 * ref) < 2e-3


@pytest.mark.parametrize("n_out,k_in,T", [(6400, 1600, 2048), (1600, 6400, 2048), (136, 72, 200), (72, 136, 200)])
def test_wgrad_orientation(n_out, k_in, T):
    """kernels.wgrad picks the wide orientation (transposed-store epilogue) for tall results."""
    k = _k()
    torch.manual_seed(5)
    dy = _bf(torch.randn(T, n_out, device="cuda"))
    x = _bf(torch.randn(T, k_in, device="cuda"))
    out = torch.full((n_out, k_in), 3.0, device="cuda")
    k.wgrad(dy, x, out=out)
    assert _rel_err(out, dy.float().t() @ x.float()) < 2e-3


def test_gemm_strided_column_slice_operand():
    """wgrad of a fused projection reads a column slice of dy (lda = full width)."""
    k = _k()
    dy = _bf(torch.randn(512, 3 * 256, device="cuda"))
    x = _bf(torch.randn(512, 320, device="cuda"))
    sl = dy[:, 256:512]
    out = k.gemm(sl, x, a_mn=True, b_mn=True, epilogue=k.EPI_F32)
    assert _rel_err(out, sl.float().t() @ x.float()) < 2e-3


def test_gemm_residual_and_swiglu_epilogues():
    k = _k()
    a = _bf(torch.randn(256, 128, device="cuda"))
    b = _bf(torch.randn(384, 128, device="cuda"))
    res = torch.randn(256, 384, device="cuda")
    out = k.gemm(a, b, epilogue=k.EPI_F32_RESID, aux=res)
    assert _rel_err(out, res + a.float() @ b.float().t()) < 2e-3
    acc = a.float() @ b.float().t()
    ref = torch.nn.functional.silu(acc[:, 0::2]) * acc[:, 1::2]
    out = k.gemm(a, b, epilogue=k.EPI_BF16_SWIGLU)
    assert out.shape == (256, 192) and _rel_err(out, ref) < 1e-2


def test_gemm_rope_epilogue_and_fused_inverse_in_fa_bwd():
    """QKV GEMM with RoPE fused in the epilogue == GEMM then rope kernel; FA backward un-rotates dq/dk."""
    k = _k()
    torch.manual_seed(3)
    B, S, H, dh = 2, 128, 3, 64
    d = H * dh
    cache = F_ref.rope_cache(256, dh, device="cuda")
    x = _bf(torch.randn(B * S, d, device="cuda"))
    w = _bf(torch.randn(3 * d, d, device="cuda") * 0.05)
    fused = k.gemm(x, w, rope=(cache, S, dh, 2 * d)).view(B, S, 3, H, dh)
    plain = k.gemm(x, w).view(B, S, 3, H, dh).clone()
    k.rope_qk_inplace(plain, cache)
    assert _rel_err(fused, plain) < 6e-3
    assert torch.equal(fused[:, :, 2], plain[:, :, 2])
    # backward: fused un-rotation vs explicit inverse-rope kernel
    q, kk, v = (fused[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    o, lse = k.fa2_fwd(q, kk, v, True)
    do = _bf(torch.randn(B, H, S, dh, device="cuda"))
    a = torch.empty(B, S, 3, H, dh, dtype=torch.bfloat16, device="cuda")
    b = torch.empty_like(a)
    k.fa2_bwd(q, kk, v, o, lse, do, True, *(a[:, :, i].permute(0, 2, 1, 3) for i in range(3)), rope_cache=cache)
    k.fa2_bwd(q, kk, v, o, lse, do, True, *(b[:, :, i].permute(0, 2, 1, 3) for i in range(3)))
    k.rope_qk_inplace(b, cache, inverse=True)
    assert _rel_err(a, b) < 1e-2


# ------------------------------------------------------------------------------------------------ norms
@pytest.mark.parametrize("T,d", [(64, 768), (2048, 1600), (33, 2560), (7, 100), (300, 3072), (16, 4096)])
def test_rmsnorm_fwd_bwd(T, d):
    k = _k()
    torch.manual_seed(0)
    x = torch.randn(T, d, device="cuda")
    w = torch.rand(d, device="cuda") + 0.5
    y, rstd, _ = k.rmsnorm_fwd(x, w, 1e-5, torch.float32)
    torch.testing.assert_close(y, F_ref.rmsnorm_ref(x, w, 1e-5), rtol=1e-4, atol=1e-5)
    yb, _, _ = k.rmsnorm_fwd(x, w, 1e-5, torch.bfloat16)
    assert _rel_err(yb, F_ref.rmsnorm_ref(x, w, 1e-5)) < 5e-3
    # fused residual
    r = _bf(torch.randn(T, d, device="cuda"))
    h, _, xo = k.rmsnorm_fwd(x, w, 1e-5, torch.bfloat16, resid=r)
    torch.testing.assert_close(xo, x + r.float())
    assert _rel_err(h, F_ref.rmsnorm_ref(x + r.float(), w, 1e-5)) < 5e-3
    # backward vs autograd
    xr = x.clone().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    dy = torch.randn(T, d, device="cuda")
    dres = torch.randn(T, d, device="cuda")
    F_ref.rmsnorm_ref(xr, wr, 1e-5).backward(dy)
    dx, dxb, dw = k.rmsnorm_bwd(dy, x, w, rstd, dres, True)
    torch.testing.assert_close(dx, xr.grad + dres, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dw, wr.grad, rtol=1e-4, atol=1e-3)
    assert _rel_err(dxb, dx) < 5e-3


def test_residual_add_and_cast():
    k = _k()
    x = torch.randn(1000, 37, device="cuda")
    y = _bf(torch.randn(1000, 37, device="cuda"))
    torch.testing.assert_close(k.residual_add(x, y), x + y.float())
    assert torch.equal(k.cast_bf16(x), x.to(torch.bfloat16))


# ------------------------------------------------------------------------------------------------ rope / swiglu / embedding
def test_rope_inplace_and_inverse():
    k = _k()
    B, S, H, dh = 2, 48, 3, 64
    cache = F_ref.rope_cache(64, dh, device="cuda")
    qkv = _bf(torch.randn(B, S, 3, H, dh, device="cuda"))
    orig = qkv.clone()
    k.rope_qk_inplace(qkv, cache)
    for i in range(2):
        ref = F_ref.rope_ref(orig[:, :, i].permute(0, 2, 1, 3).float(), cache)  # (B,H,S,dh)
        assert _rel_err(qkv[:, :, i].permute(0, 2, 1, 3), ref) < 5e-3
    assert torch.equal(qkv[:, :, 2], orig[:, :, 2])
    k.rope_qk_inplace(qkv, cache, inverse=True)
    assert _rel_err(qkv, orig) < 8e-3


def test_swiglu():
    k = _k()
    ab = _bf(torch.randn(300, 2 * 256, device="cuda"))
    a, b = ab[:, :256].float(), ab[:, 256:].float()
    assert _rel_err(k.swiglu_fwd(ab), F_ref.swiglu_ref(a, b)) < 5e-3
    ar, br = a.clone().requires_grad_(True), b.clone().requires_grad_(True)
    dg = _bf(torch.randn(300, 256, device="cuda"))
    F_ref.swiglu_ref(ar, br).backward(dg.float())
    dab = k.swiglu_bwd(ab, dg)
    assert _rel_err(dab[:, :256], ar.grad) < 6e-3 and _rel_err(dab[:, 256:], br.grad) < 6e-3


def test_embedding():
    k = _k()
    w = torch.randn(1000, 96, device="cuda")
    ids = torch.randint(0, 1000, (4, 50), device="cuda")
    torch.testing.assert_close(k.embedding_fwd(ids, w), w[ids])
    dout = torch.randn(4, 50, 96, device="cuda")
    dw = torch.full_like(w, 7.0)
    k.embedding_bwd(ids, dout, dw, zero_first=True)
    ref = torch.zeros_like(w).index_put_((ids.reshape(-1),), dout.reshape(-1, 96), accumulate=True)
    torch.testing.assert_close(dw, ref, rtol=1e-5, atol=1e-5)


# ------------------------------------------------------------------------------------------------ cross-entropy
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("T,V", [(64, 10000), (2048, 10000), (5, 33)])
def test_cross_entropy(dtype, T, V):
    k = _k()
    torch.manual_seed(0)
    logits = (torch.randn(T, V, device="cuda") * 3).to(dtype)
    tgt = torch.randint(0, V, (T,), device="cuda")
    loss, lse = k.xent_fwd(logits, tgt)
    lr = logits.float().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(lr, tgt)
    torch.testing.assert_close(loss, ref.detach(), rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(lse, torch.logsumexp(logits.float(), -1), rtol=1e-4, atol=1e-4)
    ref.backward()
    dl = k.xent_bwd(logits, tgt, lse, torch.tensor(1.0, device="cuda"))
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-4
    assert _rel_err(dl, lr.grad) < tol


# ------------------------------------------------------------------------------------------------ AdamW / clip
def test_adamw_multi_tensor_matches_course_arithmetic():
    from cs336_a2_systems_b200.optim import AdamW
    torch.manual_seed(0)
    shapes = [(1000, 33), (7,), (4096, 16), (5000,)]
    ps = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in shapes]
    qs = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    fused = AdamW(ps, lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1)
    plain = AdamW(qs, lr=1e-2, betas=(0.9, 0.95), eps=1e-8, weight_decay=0.1, fused=False)
    for step in range(4):
        for p, q in zip(ps, qs):
            g = torch.randn_like(p)
            p.grad = g.clone()
            q.grad = g.clone()
        fused.step()
        plain.step()
    for p, q in zip(ps, qs):
        torch.testing.assert_close(p, q, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(fused.state[p]["v"], plain.state[q]["v"], rtol=1e-5, atol=1e-8)
    assert fused.state[ps[0]]["t"] == 5
    fused.sync_step_from_device()
    assert fused.state[ps[0]]["t"] == 5


def test_clip_grad_norm():
    from cs336_a2_systems_b200.ops import clip_gradient
    ps = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in [(300, 70), (11,), (9000,)]]
    for p in ps:
        p.grad = torch.randn_like(p) * 3
    gs = [p.grad.clone() for p in ps]
    total = torch.sqrt(sum((g ** 2).sum() for g in gs))
    norm = clip_gradient(ps, 1.0)
    torch.testing.assert_close(norm, total, rtol=1e-4, atol=1e-4)
    coef = min(1.0, 1.0 / (float(total) + 1e-6))
    for p, g in zip(ps, gs):
        torch.testing.assert_close(p.grad, g * coef, rtol=1e-4, atol=1e-6)


# ------------------------------------------------------------------------------------------------ row softmax
@pytest.mark.parametrize("rows,cols", [(1, 1), (5, 7), (33, 1000), (4, 10000), (2, 16384), (2, 20011)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_softmax_rows(rows, cols, dtype):
    k = _k()
    torch.manual_seed(0)
    x = (torch.randn(rows, cols, device="cuda") * 3).to(dtype)
    y = k.softmax_rows(x, out_dtype=torch.float32, scale=0.7)
    ref = torch.softmax(x.float() * 0.7, -1)
    torch.testing.assert_close(y, ref, rtol=2e-5, atol=1e-7)
    torch.testing.assert_close(y.sum(-1), torch.ones(rows, device="cuda"), rtol=1e-5, atol=1e-5)
    yb = k.softmax_rows(x, scale=0.7)  # same dtype out
    assert yb.dtype == dtype and _rel_err(yb, ref) < (1e-6 if dtype == torch.float32 else 4e-3)


def test_softmax_rows_causal_and_strided():
    k = _k()
    torch.manual_seed(1)
    B, N = 3, 200
    s = torch.randn(B * N, N + 8, device="cuda").bfloat16()[:, :N]  # row stride N + 8
    p = k.softmax_rows(s, out_dtype=torch.float32, scale=0.125, causal_period=N)
    mask = torch.ones(N, N, device="cuda", dtype=torch.bool).tril().repeat(B, 1)
    ref = torch.softmax((s.float() * 0.125).masked_fill(~mask, float("-inf")), -1)
    torch.testing.assert_close(p, ref, rtol=2e-5, atol=1e-7)
    assert float(p[~mask].abs().max()) == 0.0


def test_softmax_public_op_with_autograd():
    from cs336_a2_systems_b200.ops import softmax
    torch.manual_seed(2)
    x = torch.randn(4, 6, 333, device="cuda", requires_grad=True)
    w = torch.randn(4, 6, 333, device="cuda")
    (softmax(x, -1) * w).sum().backward()
    x2 = x.detach().clone().requires_grad_(True)
    (torch.softmax(x2, -1) * w).sum().backward()
    torch.testing.assert_close(x.grad, x2.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(softmax(x.detach(), 1), torch.softmax(x.detach(), 1))  # other dims: oracle path


@pytest.mark.parametrize("N,D,causal,dtype", [(256, 64, True, torch.bfloat16), (200, 16, False, torch.bfloat16),
                                               (384, 128, True, torch.float32)])
def test_materialized_attention_matches_oracle(N, D, causal, dtype):
    """The O(N^2)-memory comparison arm (GEMM + fused mask/softmax + GEMM, and its backward) against the oracle."""
    from cs336_a2_systems_b200.ops import native
    torch.manual_seed(3)
    q, kk, v = (torch.randn(2, N, D, device="cuda").to(dtype).requires_grad_(True) for _ in range(3))
    do = torch.randn(2, N, D, device="cuda").to(dtype)
    o = native.materialized_attention(q, kk, v, causal)
    o.backward(do)
    q2, k2, v2 = (t.detach().float().requires_grad_(True) for t in (q, kk, v))
    o_ref, _ = F_ref.attention_and_lse_ref(q2, k2, v2, causal)
    o_ref.backward(do.float())
    assert _rel_err(o, o_ref) < 1.5e-2
    for a, b in ((q, q2), (kk, k2), (v, v2)):
        assert _rel_err(a.grad, b.grad) < 3e-2


# ------------------------------------------------------------------------------------------------ wgrad + AdamW epilogue
@pytest.mark.parametrize("T,rows,N", [(512, (256,), 128), (2048, (1600, 1600, 1600), 1600), (1024, (640, 640), 320),
                                      (300, (200,), 136), (2048, (1600,), 6400)])
@pytest.mark.parametrize("pair", [False, True])
def test_wgrad_adamw_epilogue(T, rows, N, pair):
    """dW = dy^T x consumed by AdamW inside the GEMM epilogue == wgrad GEMM into a gradient tensor + the multi-tensor
    AdamW kernel (same course arithmetic, optimizer.py:50-86): p, m, v to fp32 rounding, identical bf16 shadow."""
    k = _k()
    torch.manual_seed(0)
    M = sum(rows)
    if pair and N < 256:
        pytest.skip("CTA pairs need at least one 256-row tile")
    dy = _bf(torch.randn(T, M, device="cuda") * 0.1)
    x = _bf(torch.randn(T, N, device="cuda"))
    ps = [torch.randn(r, N, device="cuda") * 0.05 for r in rows]
    ms = [torch.randn(r, N, device="cuda") * 0.01 for r in rows]
    vs = [torch.rand(r, N, device="cuda") * 1e-3 for r in rows]
    # reference: gradient tensor, then the multi-tensor kernel
    g = k.wgrad(dy, x)
    p_ref, m_ref, v_ref = [t.clone() for t in ps], [t.clone() for t in ms], [t.clone() for t in vs]
    sh_ref = [torch.empty_like(t, dtype=torch.bfloat16) for t in ps]
    gs, r0 = [], 0
    for r in rows:
        gs.append(g[r0:r0 + r].contiguous())
        r0 += r
    launch = k.AdamWLaunch(p_ref, gs, m_ref, v_ref, sh_ref, step0=3, mode=0)
    launch.set_hyper(3e-3, 0.9, 0.95, 1e-8, 0.1)
    launch.launch()
    # fused
    dh = k.DeviceHyper("cuda", step0=3)
    dh.set(3e-3, 0.9, 0.95, 1e-8, 0.1)
    shadow = torch.empty(M, N, dtype=torch.bfloat16, device="cuda")
    gout = torch.empty(M, N, device="cuda")
    k.wgrad_adamw(dy, x, ps, ms, vs, shadow, dh, mode=0, grad_out=gout, pair=pair)
    torch.cuda.synchronize()
    torch.testing.assert_close(gout, g, rtol=1e-5, atol=1e-5)
    r0 = 0
    for i, r in enumerate(rows):
        torch.testing.assert_close(ms[i], m_ref[i], rtol=1e-5, atol=1e-7)
        torch.testing.assert_close(vs[i], v_ref[i], rtol=1e-5, atol=1e-9)
        torch.testing.assert_close(ps[i], p_ref[i], rtol=1e-5, atol=1e-6)
        assert _rel_err(shadow[r0:r0 + r], sh_ref[i]) < 1e-3
        r0 += r


# ------------------------------------------------------------------------------------------------ SwiGLU in the GEMM epilogues
@pytest.mark.parametrize("T,K,F", [(256, 128, 128), (2048, 1600, 6400), (512, 320, 896), (300, 256, 384)])
def test_gemm_swiglu_training_epilogues(T, K, F):
    """Forward: ONE GEMM over [w1; w3] whose epilogue writes silu(a)*b and keeps a | b; backward: the dgrad GEMM of the
    down-projection applies the gate's backward in its epilogue.  Against fp32 PyTorch."""
    k = _k()
    torch.manual_seed(0)
    x = _bf(torch.randn(T, K, device="cuda"))
    w13 = _bf(torch.randn(2 * F, K, device="cuda") / math.sqrt(K))
    assert k.swiglu_up_supported(x, w13)
    g, ab = k.gemm_swiglu_fwd(x, w13)
    ref_ab = x.float() @ w13.float().t()
    a, b = ref_ab[:, :F], ref_ab[:, F:]
    assert _rel_err(ab, ref_ab) < 6e-3
    assert _rel_err(g, torch.nn.functional.silu(a) * b) < 8e-3
    # backward epilogue, against autograd through the bf16-rounded pre-activations the kernel reads
    d = 192
    dy = _bf(torch.randn(T, d, device="cuda"))
    w2 = _bf(torch.randn(d, F, device="cuda") / math.sqrt(d))
    dab = k.gemm_swiglu_bwd(dy, w2, ab)
    abf = ab.float().requires_grad_(True)
    (torch.nn.functional.silu(abf[:, :F]) * abf[:, F:]).backward(dy.float() @ w2.float())
    assert _rel_err(dab, abf.grad) < 8e-3
