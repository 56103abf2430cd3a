"""Hardware check of the opt-in FlashAttention kernel variants against a materialised fp32 oracle, plus timings on
the 16x16384x64 causal shape.  The variants are selected by environment variables that the library reads once per
process, so set them on the command line:

    python tools/fa_variants_check.py                                  # default kernels
    B200SYS_FA_FWD_WG2=1 python tools/fa_variants_check.py             # forward, two softmax warpgroups (2: + O in TMEM)
    B200SYS_FA_BWD_WG=2 python tools/fa_variants_check.py              # backward, 2 (or 4) softmax column groups
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from cs336_a2_systems_b200.ops import kernels as k  # noqa: E402

print("variants:", {v: os.environ.get(v, "-") for v in ("B200SYS_FA_FWD_WG2", "B200SYS_FA_BWD_WG")}, flush=True)
torch.manual_seed(0)
ok = True
for (B, H, N, D, causal) in [(2, 3, 512, 64, True), (1, 2, 1000, 64, False), (1, 2, 384, 128, True), (1, 1, 300, 16, True),
                             (1, 2, 640, 80, True)]:
    q, kk, v = (torch.randn(B, H, N, D, device="cuda").bfloat16().requires_grad_(True) for _ in range(3))
    do = torch.randn(B, H, N, D, device="cuda").bfloat16()
    o, lse = k.fa2_fwd(q.detach(), kk.detach(), v.detach(), causal)
    dq, dk, dv = k.fa2_bwd(q.detach(), kk.detach(), v.detach(), o, lse, do, causal)
    s = (q.float() @ kk.float().transpose(-1, -2)) / D ** 0.5
    if causal:
        s = s.masked_fill(torch.triu(torch.ones(N, N, device="cuda", dtype=torch.bool), 1), float("-inf"))
    ref = torch.softmax(s, -1) @ v.float()
    ref.backward(do.float())
    errs = {"O": (o.float() - ref).abs().max().item(), "L": (lse - torch.logsumexp(s, -1)).abs().max().item()}
    for name, got, want in (("dQ", dq, q.grad), ("dK", dk, kk.grad), ("dV", dv, v.grad)):
        errs[name] = ((got.float() - want).norm() / (want.norm() + 1e-12)).item()
    print(f"B{B} H{H} N{N} D{D} causal={causal}: " + " ".join(f"{a}={b:.2e}" for a, b in errs.items()), flush=True)
    ok &= errs["O"] < 3e-2 and errs["L"] < 2e-3 and max(errs["dQ"], errs["dK"], errs["dV"]) < 2e-2


def timed(fn, n=10):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


for (B, H, N, D) in [(1, 16, 16384, 64), (100, 1, 512, 64), (1, 8, 16384, 128)]:
    q, kk, v, do = (torch.randn(B, H, N, D, device="cuda").bfloat16() for _ in range(4))
    o, lse = k.fa2_fwd(q, kk, v, True)
    print(f"{B}x{H}x{N}x{D} causal: fwd {timed(lambda: k.fa2_fwd(q, kk, v, True)):.3f} ms, "
          f"bwd {timed(lambda: k.fa2_bwd(q, kk, v, o, lse, do, True)):.3f} ms", flush=True)
print("PARITY_OK" if ok else "PARITY_FAIL")
