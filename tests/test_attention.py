"""FlashAttention-2 Functions vs. a materialised-softmax oracle (reference ``tests/test_attention.py:11-104``:
same inputs B=4, N=128, D=64, seed 0, same oracle, same 1e-2 tolerances, logsumexp found among the saved tensors)."""
import pytest
import torch

from cs336_a2_systems_b200.ops.functional import attention_and_lse_ref

from .adapters import (get_flashattention_autograd_function_pytorch,
                       get_flashattention_autograd_function_triton)


def _inputs(device, B=4, N=128, D=64, dtype=torch.float32, nk=None):
    g = torch.Generator().manual_seed(0)
    nk = nk or N
    q = torch.randn(B, N, D, generator=g).to(device=device, dtype=dtype).requires_grad_(True)
    k = torch.randn(B, nk, D, generator=g).to(device=device, dtype=dtype).requires_grad_(True)
    v = torch.randn(B, nk, D, generator=g).to(device=device, dtype=dtype).requires_grad_(True)
    do = torch.randn(B, N, D, generator=g).to(device=device, dtype=dtype)
    return q, k, v, do


def _check_forward(impl, device, is_causal, tol=1e-2, **kw):
    q, k, v, _ = _inputs(device, **kw)
    o = impl(q, k, v, is_causal)
    o_ref, l_ref = attention_and_lse_ref(q, k, v, is_causal)
    saved = [t for t in o.grad_fn.saved_tensors if t.shape == (q.shape[0], q.shape[1])]
    assert len(saved) == 1, "exactly one saved tensor must have the logsumexp's shape (B, N_q)"
    torch.testing.assert_close(o.float(), o_ref, rtol=tol, atol=tol)
    torch.testing.assert_close(saved[0].float(), l_ref, rtol=tol, atol=tol)


def _check_backward(impl, device, is_causal, tol=1e-2, **kw):
    q, k, v, do = _inputs(device, **kw)
    impl(q, k, v, is_causal).backward(do)
    q2, k2, v2 = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
    attention_and_lse_ref(q2, k2, v2, is_causal)[0].backward(do.float())
    for a, b in ((q, q2), (k, k2), (v, v2)):
        torch.testing.assert_close(a.grad.float(), b.grad.float(), rtol=tol, atol=tol)


@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_forward_pass_pytorch(is_causal):
    _check_forward(get_flashattention_autograd_function_pytorch().apply, "cpu", is_causal)


@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_backward_pytorch(is_causal):
    _check_backward(get_flashattention_autograd_function_pytorch().apply, "cpu", is_causal)


def test_flash_pytorch_ragged_and_2d():
    """Sequence lengths that are not tile multiples, Nq != Nk, and un-batched (N, D) inputs."""
    f = get_flashattention_autograd_function_pytorch().apply
    _check_forward(f, "cpu", False, B=2, N=100, D=16, nk=75)
    _check_backward(f, "cpu", False, B=2, N=100, D=16, nk=75)
    q, k, v, _ = _inputs("cpu", B=1, N=48, D=32)
    o = f(q[0], k[0], v[0], True)
    o_ref, _ = attention_and_lse_ref(q, k, v, True)
    torch.testing.assert_close(o, o_ref[0], rtol=1e-4, atol=1e-4)


def test_flash_pytorch_random_shapes():
    """Property test: the tiled PyTorch FlashAttention-2 (forward output, saved logsumexp, dQ/dK/dV) agrees with the
    materialised oracle for arbitrary small shapes — lengths that straddle tile boundaries, Nq != Nk, causal or not."""
    from hypothesis import given, settings, strategies as st

    f = get_flashattention_autograd_function_pytorch().apply

    @settings(max_examples=25, deadline=None)
    @given(B=st.integers(1, 3), N=st.integers(1, 70), nk=st.integers(1, 70), D=st.sampled_from([4, 8, 24]),
           causal=st.booleans(), seed=st.integers(0, 2 ** 16))
    def check(B, N, nk, D, causal, seed):
        if causal:
            nk = N  # the causal mask is defined for aligned self-attention
        g = torch.Generator().manual_seed(seed)
        q, k, v = (torch.randn(B, n, D, generator=g, requires_grad=True) for n in (N, nk, nk))
        do = torch.randn(B, N, D, generator=g)
        o = f(q, k, v, causal)
        lse = [t.clone() for t in o.grad_fn.saved_tensors if t.shape == (B, N)]
        if D != N:  # with D == N the saved O has the same shape as the logsumexp
            assert len(lse) == 1
        o.backward(do)
        q2, k2, v2 = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
        o_ref, l_ref = attention_and_lse_ref(q2, k2, v2, causal)
        o_ref.backward(do)
        torch.testing.assert_close(o, o_ref, rtol=1e-4, atol=1e-4)
        assert any(torch.allclose(t, l_ref, rtol=1e-4, atol=1e-4) for t in lse)
        for a, b in ((q, q2), (k, k2), (v, v2)):
            torch.testing.assert_close(a.grad, b.grad, rtol=1e-3, atol=1e-4)

    check()


def test_lazy_rescale_recurrence_is_exact():
    """The recurrence of the forward kernel (``fa2_fwd_kernel``: O resident in tensor memory), restated in PyTorch: the
    exponentials are taken against a reference maximum that only moves when the true maximum exceeds it by 2^tau, O and
    l are rescaled only then, and L = m_ref·ln2 + ln(l).  Whatever the reference, O / l and L must equal the exact
    softmax attention; P never exceeds 2^tau."""
    import math

    torch.manual_seed(3)
    B, N, D, T, tau = 2, 96, 16, 16, 8.0
    q, k, v = torch.randn(B, N, D) * 3, torch.randn(B, N, D) * 3, torch.randn(B, N, D)
    scale_log2 = (1.0 / math.sqrt(D)) * math.log2(math.e)
    o_ref, l_ref = attention_and_lse_ref(q, k, v, True)
    m_ref = torch.full((B, N), -float("inf"))
    l = torch.zeros(B, N)
    o = torch.zeros(B, N, D)
    p_max, n_moves = 0.0, 0
    rows = torch.arange(N)[:, None]
    for j in range(N // T):
        s = torch.einsum("bqd,bkd->bqk", q, k[:, j * T:(j + 1) * T]) * scale_log2
        mask = rows >= (torch.arange(T)[None, :] + j * T)
        s = torch.where(mask, s, torch.full_like(s, -float("inf")))
        m_tile = s.amax(-1)
        move = m_tile > m_ref + tau
        m_new = torch.where(move, m_tile, m_ref)
        m_use = torch.where(torch.isinf(m_new), torch.zeros_like(m_new), m_new)
        alpha = torch.where(move, torch.exp2(m_ref - m_use), torch.ones_like(m_ref))
        p = torch.where(mask, torch.exp2(s - m_use[..., None]), torch.zeros_like(s))
        p_max = max(p_max, float(p.max()))
        n_moves += int(move.sum())
        o = o * alpha[..., None] + torch.einsum("bqk,bkd->bqd", p, v[:, j * T:(j + 1) * T])
        l = l * alpha + p.sum(-1)
        m_ref = m_new
    torch.testing.assert_close(o / l[..., None], o_ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close((m_ref + torch.log2(l)) * math.log(2.0), l_ref, rtol=1e-5, atol=1e-5)
    assert p_max <= 2.0 ** tau * (1 + 1e-6)
    assert n_moves < B * N * (N // T) // 2  # the rescale is the exception, not the rule


@pytest.mark.gpu
@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_forward_pass_sm100(is_causal):
    _check_forward(get_flashattention_autograd_function_triton().apply, "cuda", is_causal)


@pytest.mark.gpu
@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_backward_sm100(is_causal):
    _check_backward(get_flashattention_autograd_function_triton().apply, "cuda", is_causal)


@pytest.mark.gpu
@pytest.mark.parametrize("N,D", [(128, 16), (256, 32), (512, 64), (1024, 128), (384, 80), (200, 64)])
@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_sm100_shapes_bf16(N, D, is_causal):
    f = get_flashattention_autograd_function_triton().apply
    _check_forward(f, "cuda", is_causal, tol=2e-2, B=2, N=N, D=D, dtype=torch.bfloat16)
    _check_backward(f, "cuda", is_causal, tol=3e-2, B=2, N=N, D=D, dtype=torch.bfloat16)


@pytest.mark.gpu
@pytest.mark.parametrize("N,D,B", [(4096, 64, 2), (4096, 128, 1), (16384, 64, 1), (2176, 80, 2)])
def test_flash_sm100_long_sequences(N, D, B):
    """Long causal sequences (many key tiles per query tile, cluster-shared K/V) against the materialised oracle."""
    f = get_flashattention_autograd_function_triton().apply
    _check_forward(f, "cuda", True, tol=2e-2, B=B, N=N, D=D, dtype=torch.bfloat16)
    _check_backward(f, "cuda", True, tol=3e-2, B=B, N=N, D=D, dtype=torch.bfloat16)


@pytest.mark.gpu
@pytest.mark.parametrize("N,nk,D", [(256, 640, 64), (640, 256, 128), (300, 1000, 80), (1000, 200, 16)])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_flash_sm100_cross_lengths(N, nk, D, dtype):
    """N_q != N_k (non-causal; the causal mask is defined for aligned self-attention), ragged tails, d = 80 / 128."""
    f = get_flashattention_autograd_function_triton().apply
    _check_forward(f, "cuda", False, tol=2e-2, B=2, N=N, D=D, nk=nk, dtype=dtype)
    _check_backward(f, "cuda", False, tol=3e-2, B=2, N=N, D=D, nk=nk, dtype=dtype)


@pytest.mark.gpu
@pytest.mark.parametrize("N,D", [(512, 64), (384, 128), (200, 80)])
@pytest.mark.parametrize("is_causal", [False, True])
def test_flash_sm100_fp32_path_has_no_torch_casts(N, D, is_causal):
    """fp32 callers: ONE fused cast launch for q, k, v on the way in, fp32 O and fp32 dQ/dK/dV written by the kernels
    (the saved O is the fp32 tensor the caller got, not a bf16 rounding of it)."""
    f = get_flashattention_autograd_function_triton().apply
    q, k, v, do = _inputs("cuda", B=2, N=N, D=D)
    o = f(q, k, v, is_causal)
    assert o.dtype == torch.float32
    saved_o = [t for t in o.grad_fn.saved_tensors if t.shape == o.shape and t.dtype == torch.float32]
    assert len(saved_o) == 1 and saved_o[0].data_ptr() == o.data_ptr()
    o.backward(do)
    assert q.grad.dtype == k.grad.dtype == v.grad.dtype == torch.float32
    o_ref, _ = attention_and_lse_ref(q, k, v, is_causal)
    torch.testing.assert_close(o, o_ref, rtol=1e-2, atol=1e-2)


@pytest.mark.gpu
def test_flash_sm100_cluster_and_plain_launch_agree_bitwise():
    """The 2-CTA cluster launch (K/V tiles multicast to both CTAs) runs the same arithmetic as the plain launch."""
    from cs336_a2_systems_b200.ops import kernels
    torch.manual_seed(5)
    for (B, H, N, D, causal) in [(2, 3, 1024, 64, True), (1, 2, 768, 128, True), (1, 2, 512, 80, False)]:
        q, k, v = (torch.randn(B, H, N, D, device="cuda").bfloat16() for _ in range(3))
        o1, l1 = kernels.fa2_fwd(q, k, v, causal, cluster=True)
        o0, l0 = kernels.fa2_fwd(q, k, v, causal, cluster=False)
        assert torch.equal(o1, o0) and torch.equal(l1, l0)


@pytest.mark.gpu
def test_flash_sm100_backward_run_to_run():
    """dK and dV are accumulated in tensor memory by one CTA per key tile: bit-identical from run to run.  dQ is summed
    across key tiles with fp32 ``red.global.add`` in whatever order the CTAs arrive: equal to fp32 rounding only."""
    from cs336_a2_systems_b200.ops import kernels
    torch.manual_seed(6)
    B, H, N, D = 1, 4, 2048, 64
    q, k, v, do = (torch.randn(B, H, N, D, device="cuda").bfloat16() for _ in range(4))
    o, lse = kernels.fa2_fwd(q, k, v, True)
    runs = [kernels.fa2_bwd(q, k, v, o, lse, do, True, grads_f32=True) for _ in range(3)]
    torch.cuda.synchronize()
    for dq, dk, dv in runs[1:]:
        assert torch.equal(dk, runs[0][1]) and torch.equal(dv, runs[0][2])
        rel = (dq - runs[0][0]).abs().max() / runs[0][0].abs().max()
        assert float(rel) < 1e-5, float(rel)
