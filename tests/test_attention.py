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
    """The recurrence of the O-in-tensor-memory forward variant (``fa2_fwd_wg2o_kernel``), restated in PyTorch: the
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
@pytest.mark.skipif(__import__("os").environ.get("B200SYS_TEST_EXPERIMENTAL") != "1",
                    reason="opt-in kernel variants, validated on a few shapes only (set B200SYS_TEST_EXPERIMENTAL=1)")
@pytest.mark.parametrize("env", [{"B200SYS_FA_FWD_WG2": "1"}, {"B200SYS_FA_FWD_WG2": "2"}, {"B200SYS_FA_FWD_WG2": "3"},
                                 {"B200SYS_FA_BWD_WG": "2"},
                                 {"B200SYS_FA_BWD_WG": "4"}, {"B200SYS_FA_BWD_WG": "2r"}, {"B200SYS_FA_BWD_WG": "4r"},
                                 {"B200SYS_FA_FWD_WG2": "1", "B200SYS_FA_BWD_WG": "2"}])
def test_flash_kernel_variants_match_oracle(env):
    """The opt-in FlashAttention variants (two softmax warpgroups in the forward, 2 / 4 softmax column groups in the
    backward) are selected by environment variables that the library reads once per process, hence the subprocess
    running ``tools/fa_variants_check.py`` (parity of O, L, dQ, dK, dV with a materialised fp32 oracle)."""
    import os
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "tools", "fa_variants_check.py")], env=dict(os.environ, **env),
                         capture_output=True, text=True, timeout=300, cwd=root)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "PARITY_OK" in out.stdout, out.stdout[-2000:]
