"""Hardware check of the FlashAttention kernels against a materialised fp32 oracle on awkward shapes (ragged tails,
head dims 16 / 80 / 128, causal or not), plus timings on the leaderboard shape (16 x 16384 x 64 causal), the LM shape
(100 x 512 x 64) and a head-dim-128 shape, each with the 2-CTA cluster launch (K/V multicast) and the plain launch.

    python tools/fa_check.py [--json out.json]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from cs336_a2_systems_b200.ops import kernels as k  # noqa: E402
from cs336_a2_systems_b200.utils.clocks import ClockSampler  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--json", default=None)
args = ap.parse_args()
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


def timed(fn, n=20):
    for _ in range(3):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


rows = []
for (B, H, N, D) in [(1, 16, 16384, 64), (100, 1, 512, 64), (4, 25, 512, 64), (1, 8, 16384, 128), (1, 32, 4096, 80)]:
    q, kk, v, do = (torch.randn(B, H, N, D, device="cuda").bfloat16() for _ in range(4))
    o, lse = k.fa2_fwd(q, kk, v, True)
    with ClockSampler(0, period_ms=50) as clk:
        f1 = timed(lambda: k.fa2_fwd(q, kk, v, True, cluster=True))
        f0 = timed(lambda: k.fa2_fwd(q, kk, v, True, cluster=False))
        bw = timed(lambda: k.fa2_bwd(q, kk, v, o, lse, do, True))
    fl = 4.0 * B * H * N * N * D * 0.5
    row = dict(B=B, H=H, N=N, D=D, causal=True, fwd_cluster_ms=f1, fwd_plain_ms=f0, bwd_ms=bw,
               fwd_tflops=fl / min(f0, f1) / 1e9, bwd_tflops=2.5 * fl / bw / 1e9, clocks=clk.summary())
    rows.append(row)
    print(f"{B}x{H}x{N}x{D} causal: fwd cluster {f1:.3f} ms / plain {f0:.3f} ms ({row['fwd_tflops']:.0f} TF/s), "
          f"bwd {bw:.3f} ms ({row['bwd_tflops']:.0f} TF/s) clocks {row['clocks'].get('sm_mhz')} MHz "
          f"{row['clocks'].get('reasons')}", flush=True)
if args.json:
    json.dump(rows, open(args.json, "w"), indent=1)
print("PARITY_OK" if ok else "PARITY_FAIL")
