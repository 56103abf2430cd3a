"""The allocation-free RMSNorm-backward entry point, driven through the C ABI.

Some users of the library cannot let it allocate: their process gave all device memory to its own
pool, or sits in a loop where cudaMalloc's device-wide synchronisation is not acceptable, or talks to
us through an interface that passes input and output buffers only.  For them
``b200_rmsnorm_bwd_noalloc`` keeps its partial-sum scratch in a fixed array inside the library.  This
test calls it with raw pointers (ctypes, no Python wrapper in between) and checks that it returns what
the ordinary entry point returns with caller-provided scratch — twice in a row with different shapes,
because the built-in array is shared by all calls and keeps the previous call's contents.  It also
checks that a problem too large for the array is refused with a message rather than truncated.

The file name sorts last on purpose: a ``pytest -x`` run has executed every other test before it.
"""
import ctypes
from ctypes import c_int, c_longlong, c_void_p

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return c_void_p(0 if t is None else t.data_ptr())


def _run_pair(lib, T, d, seed, with_resid):
    torch.manual_seed(seed)
    dev = "cuda"
    x = torch.randn(T, d, device=dev)
    w = torch.rand(d, device=dev) + 0.5
    dy = torch.randn(T, d, device=dev)
    dres = torch.randn(T, d, device=dev) if with_resid else None
    rstd = torch.rsqrt(x.pow(2).mean(-1) + 1e-5).contiguous()
    stream = c_void_p(torch.cuda.current_stream().cuda_stream)
    F32 = 1

    rows = lib.b200_rmsnorm_bwd_workspace_rows(c_int(T))
    scratch = torch.empty(rows * d, device=dev)
    dx_a, dw_a = torch.empty_like(x), torch.empty_like(w)
    dxb_a = torch.empty(T, d, dtype=torch.bfloat16, device=dev)
    rc = lib.b200_rmsnorm_bwd(_ptr(dy), c_int(F32), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx_a), _ptr(dxb_a),
                              _ptr(dw_a), c_int(0), _ptr(scratch), c_int(T), c_int(d), stream)
    assert rc == 0, lib.b200_last_error()

    dx_b, dw_b = torch.empty_like(x), torch.empty_like(w)
    dxb_b = torch.empty(T, d, dtype=torch.bfloat16, device=dev)
    rc = lib.b200_rmsnorm_bwd_noalloc(_ptr(dy), c_int(F32), _ptr(x), _ptr(w), _ptr(rstd), _ptr(dres), _ptr(dx_b),
                                      _ptr(dxb_b), _ptr(dw_b), c_int(0), c_int(T), c_int(d), stream)
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()
    # same tolerances as tests/test_kernels_gpu.py::test_rmsnorm_fwd_bwd
    torch.testing.assert_close(dx_b, dx_a, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dw_b, dw_a, rtol=1e-4, atol=1e-3)
    torch.testing.assert_close(dxb_b.float(), dxb_a.float(), rtol=1e-2, atol=1e-2)
    # and against plain PyTorch autograd, so "both wrong the same way" cannot pass
    xr, wr = x.clone().requires_grad_(True), w.clone().requires_grad_(True)
    y = wr * (xr * torch.rsqrt(xr.pow(2).mean(-1, keepdim=True) + 1e-5))
    y.backward(dy)
    ref_dx = xr.grad + (dres if dres is not None else 0)
    torch.testing.assert_close(dx_b, ref_dx, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dw_b, wr.grad, rtol=1e-4, atol=1e-3)


def test_rmsnorm_bwd_noalloc_matches_regular_entry_point_twice():
    from cs336_a2_systems_b200 import _native

    lib = _native.lib()
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_rmsnorm_bwd_noalloc_max_floats.restype = c_longlong
    cap = lib.b200_rmsnorm_bwd_noalloc_max_floats()
    assert cap == 304 * 4096
    # two calls in a row: the second is smaller, so it runs over the first call's leftovers
    _run_pair(lib, T=2048, d=1600, seed=0, with_resid=True)
    _run_pair(lib, T=96, d=768, seed=1, with_resid=False)
    # a third, again larger, for good measure
    _run_pair(lib, T=1024, d=2560, seed=2, with_resid=True)


def test_rmsnorm_bwd_noalloc_rejects_oversize():
    from cs336_a2_systems_b200 import _native

    lib = _native.lib()
    lib.b200_last_error.restype = ctypes.c_char_p
    T, d = 4096, 8192  # 296 rows x 8192 floats > 304 x 4096
    rc = lib.b200_rmsnorm_bwd_noalloc(c_void_p(0), c_int(1), c_void_p(0), c_void_p(0), c_void_p(0), c_void_p(0),
                                      c_void_p(0), c_void_p(0), c_void_p(0), c_int(0), c_int(T), c_int(d), c_void_p(0))
    assert rc != 0
    msg = lib.b200_last_error().decode()
    assert "scratch" in msg and "noalloc" in msg, msg
