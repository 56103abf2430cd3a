"""Launch each hot kernel at the XL-LM shapes a couple of times — the target program for ncu.
    ncu --set full --import-source on -k regex:<name> -o gpurun_out/prof python tools/ncu_targets.py [names...]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cs336_a2_systems_b200.ops import kernels as k, functional as F_ref

which = set(sys.argv[1:]) or {"gemm", "pair", "fa", "norm", "swiglu", "adamw"}
dev = "cuda"
T, d, dff, H, dh, B, S = 2048, 1600, 6400, 25, 64, 4, 512
bf = lambda *s: torch.randn(*s, device=dev).bfloat16()
reps = int(os.environ.get("NCU_REPS", "2"))
if "gemm" in which:
    x, w = bf(T, d), bf(3 * d, d) * 0.05
    for _ in range(reps):
        k.gemm(x, w)                                             # fwd NT  2048x4800x1600
    dy, w13 = bf(T, 2 * dff), bf(2 * dff, d)
    for _ in range(reps):
        k.gemm(dy, w13, b_mn=True)                               # dgrad NN 2048x1600x12800
    dy1, x1, out = bf(T, d), bf(T, d), torch.empty(d, d, device=dev)
    for _ in range(reps):
        k.gemm(dy1, x1, a_mn=True, b_mn=True, epilogue=k.EPI_F32, out=out)   # wgrad TN 1600x1600x2048
    dy2, out2 = bf(T, dff), torch.empty(dff, d, device=dev)
    for _ in range(reps):
        k.gemm(dy2, x1, a_mn=True, b_mn=True, epilogue=k.EPI_F32, out=out2)  # wgrad TN 6400x1600x2048
if "pair" in which:
    # CTA-pair (cta_group::2) kernel, forced: fwd NT 2048x12800x1600 (w1|w3) and dgrad NN 2048x1600x12800
    x, w13 = bf(T, d), bf(2 * dff, d) * 0.05
    for _ in range(reps):
        k.gemm(x, w13, pair=True)
    dy = bf(T, 2 * dff)
    for _ in range(reps):
        k.gemm(dy, w13, b_mn=True, pair=True)
if "xent" in which:
    logits, tgt = bf(T, 10000), torch.randint(0, 10000, (T,), device=dev)
    for _ in range(reps):
        loss, lse = k.xent_fwd(logits, tgt)
        k.xent_bwd(logits, tgt, lse, torch.ones((), device=dev))
if "fa" in which:
    qkv = bf(B, S, 3, H, dh)
    q, kk, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    o = torch.empty(B, S, H, dh, device=dev, dtype=torch.bfloat16).permute(0, 2, 1, 3)
    do = bf(B, S, H, dh).permute(0, 2, 1, 3)
    dqkv = torch.empty_like(qkv)
    for _ in range(reps):
        _, lse = k.fa2_fwd(q, kk, v, True, out=o)
        k.fa2_bwd(q, kk, v, o, lse, do, True, *(dqkv[:, :, i].permute(0, 2, 1, 3) for i in range(3)))
    # long-sequence benchmark shape (16 heads x 16384 x 64, causal)
    if "falong" in which:
        q2, k2, v2 = bf(1, 16, 16384, 64), bf(1, 16, 16384, 64), bf(1, 16, 16384, 64)
        o2, l2 = k.fa2_fwd(q2, k2, v2, True)
        k.fa2_bwd(q2, k2, v2, o2, l2, bf(1, 16, 16384, 64), True)
if "norm" in which:
    xx, ww, r = torch.randn(T, d, device=dev), torch.ones(d, device=dev), bf(T, d)
    for _ in range(reps):
        h, rstd, xo = k.rmsnorm_fwd(xx, ww, 1e-5, torch.bfloat16, resid=r)
        k.rmsnorm_bwd(bf(T, d), xo, ww, rstd, torch.randn(T, d, device=dev), True)
if "swiglu" in which:
    ab = bf(T, 2 * dff)
    for _ in range(reps):
        g = k.swiglu_fwd(ab)
        k.swiglu_bwd(ab, g)
if "adamw" in which:
    from cs336_a2_systems_b200.optim import AdamW
    ps = [torch.nn.Parameter(torch.randn(6400, 1600, device=dev)) for _ in range(20)]
    opt = AdamW(ps, lr=1e-3)
    for _ in range(reps):
        for p in ps:
            p.grad = torch.ones_like(p)
        opt.step()
torch.cuda.synchronize()
print("ncu targets done")
