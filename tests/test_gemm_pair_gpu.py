"""CTA-pair (tcgen05 cta_group::2) GEMM kernel vs PyTorch fp32 and vs the 1-CTA kernel."""
import pytest
import torch

pytestmark = pytest.mark.gpu

from cs336_a2_systems_b200.ops import functional as F_ref  # noqa: E402


def _k():
    from cs336_a2_systems_b200.ops import kernels
    return kernels


def _bf(t):
    return t.to(torch.bfloat16)


def _rel(a, b):
    return float((a.float() - b.float()).norm() / (b.float().norm() + 1e-12))


@pytest.mark.parametrize("M,N,K", [(256, 256, 64), (2048, 4800, 1600), (2048, 1600, 1600), (2048, 10000, 1600),
                                   (384, 200, 72), (2048, 12800, 1600)])
@pytest.mark.parametrize("bn", [0, 128, 192, 256])
def test_pair_gemm_nt(M, N, K, bn):
    k = _k()
    torch.manual_seed(0)
    a, b = _bf(torch.randn(M, K, device="cuda")), _bf(torch.randn(N, K, device="cuda"))
    c = k.gemm(a, b, force_bn=bn, pair=True)
    assert _rel(c, a.float() @ b.float().t()) < 6e-3
    c32 = k.gemm(a, b, epilogue=k.EPI_F32, force_bn=bn, pair=True)
    assert _rel(c32, a.float() @ b.float().t()) < 2e-3


@pytest.mark.parametrize("M,N,K", [(2048, 1600, 4800), (2048, 1600, 12800), (512, 256, 320)])
@pytest.mark.parametrize("bn", [0, 128, 256])
def test_pair_gemm_nn(M, N, K, bn):
    k = _k()
    torch.manual_seed(1)
    dy, w = _bf(torch.randn(M, K, device="cuda")), _bf(torch.randn(K, N, device="cuda"))
    assert _rel(k.gemm(dy, w, b_mn=True, force_bn=bn, pair=True), dy.float() @ w.float()) < 6e-3


@pytest.mark.parametrize("n_out,k_in,T", [(1600, 1600, 2048), (6400, 1600, 2048), (1600, 6400, 2048), (4800, 1600, 2048)])
def test_pair_gemm_tn_both_orientations(n_out, k_in, T):
    k = _k()
    torch.manual_seed(2)
    dy, x = _bf(torch.randn(T, n_out, device="cuda")), _bf(torch.randn(T, k_in, device="cuda"))
    ref = dy.float().t() @ x.float()
    out = k.gemm(dy, x, a_mn=True, b_mn=True, epilogue=k.EPI_F32, pair=True)
    assert _rel(out, ref) < 2e-3
    out_t = torch.empty(n_out, k_in, device="cuda")
    k.gemm(x, dy, a_mn=True, b_mn=True, epilogue=k.EPI_F32_T, out=out_t, pair=True)
    assert _rel(out_t, ref) < 2e-3


def test_pair_gemm_rope_epilogue():
    k = _k()
    B, S, H, dh = 2, 256, 4, 64
    d = H * dh
    cache = F_ref.rope_cache(256, dh, device="cuda")
    x, w = _bf(torch.randn(B * S, d, device="cuda")), _bf(torch.randn(3 * d, d, device="cuda") * 0.05)
    a = k.gemm(x, w, rope=(cache, S, dh, 2 * d), pair=True)
    b = k.gemm(x, w, rope=(cache, S, dh, 2 * d), pair=False)
    assert _rel(a, b) < 1e-3
