"""Quick hardware check of the experimental two-warpgroup FlashAttention forward (B200SYS_FA_FWD_WG2=1):
parity against a materialised fp32 oracle on a few shapes and a forward timing on the 16x16384x64 causal shape."""
import os
import sys

os.environ["B200SYS_FA_FWD_WG2"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from cs336_a2_systems_b200.ops import kernels as k  # noqa: E402

torch.manual_seed(0)
ok = True
for (B, H, N, D, causal) in [(2, 3, 512, 64, True), (1, 2, 1000, 64, False), (1, 2, 384, 128, True), (1, 1, 300, 16, True)]:
    q, kk, v = (torch.randn(B, H, N, D, device="cuda").bfloat16() for _ in range(3))
    o, lse = k.fa2_fwd(q, kk, v, causal)
    s = (q.float() @ kk.float().transpose(-1, -2)) / D ** 0.5
    if causal:
        s = s.masked_fill(torch.triu(torch.ones(N, N, device="cuda", dtype=torch.bool), 1), float("-inf"))
    ref = torch.softmax(s, -1) @ v.float()
    ref_lse = torch.logsumexp(s, -1)
    eo = (o.float() - ref).abs().max().item()
    el = (lse - ref_lse).abs().max().item()
    print(f"B{B} H{H} N{N} D{D} causal={causal}: max|dO|={eo:.3e} max|dL|={el:.3e}", flush=True)
    ok &= eo < 3e-2 and el < 2e-3
q, kk, v = (torch.randn(1, 16, 16384, 64, device="cuda").bfloat16() for _ in range(3))
for _ in range(3):
    k.fa2_fwd(q, kk, v, True)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    k.fa2_fwd(q, kk, v, True)
e1.record()
torch.cuda.synchronize()
print(f"wg2 fwd 16x16384x64 causal: {e0.elapsed_time(e1) / 10:.3f} ms (default kernel: 1.425 ms)", "PARITY_OK" if ok else "PARITY_FAIL")
