"""Launch each hot kernel at the XL-LM shapes a couple of times — the target program for ncu.
    ncu --set full --import-source on -k regex:<name> -o gpurun_out/prof python tools/ncu_targets.py [names...]
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cs336_a2_systems_b200.ops import kernels as k, functional as F_ref

which = set(sys.argv[1:]) or {"gemm", "pair", "fa", "falong", "fa128", "norm", "swiglu", "adamw", "xent", "elem", "softmax", "comm"}
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
        k.wgrad(dy1, x1, out=out)                                # wgrad TN 1600x1600x2048 (transposed-store epilogue)
    dy2, out2 = bf(T, 2 * dff), torch.empty(2 * dff, d, device=dev)
    for _ in range(reps):
        k.wgrad(dy2, x1, out=out2)                               # wgrad TN 12800x1600x2048
    # optimizer in the wgrad epilogue (EPI_ADAMW)
    pm = [torch.randn(2 * dff, d, device=dev) * 0.02, torch.zeros(2 * dff, d, device=dev), torch.zeros(2 * dff, d, device=dev)]
    dhy = k.DeviceHyper(dev, 1); dhy.set(1e-4, 0.9, 0.95, 1e-8, 0.1)
    sh = torch.empty(2 * dff, d, device=dev, dtype=torch.bfloat16)
    for _ in range(reps):
        k.wgrad_adamw(dy2, x1, [pm[0]], [pm[1]], [pm[2]], sh, dhy)
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
    # the gate inside the GEMM epilogues (training path): up-projection forward, down-projection dgrad
    xs, w13s, w2s, dys = bf(T, d), bf(2 * dff, d) * 0.03, bf(d, dff) * 0.03, bf(T, d)
    for _ in range(reps):
        g2, ab2 = k.gemm_swiglu_fwd(xs, w13s)
        k.gemm_swiglu_bwd(dys, w2s, ab2)
if "fa128" in which:
    q3, k3, v3, do3 = (bf(1, 8, 4096, 128) for _ in range(4))
    for _ in range(reps):
        o3, l3 = k.fa2_fwd(q3, k3, v3, True)
        k.fa2_bwd(q3, k3, v3, o3, l3, do3, True)
if "softmax" in which:
    sm = bf(8192, 8192)
    for _ in range(reps):
        pr = k.softmax_rows(sm, scale=0.125, causal_period=8192)
        k.softmax_bwd_rows(pr, sm, scale=0.125)
if "elem" in which:
    qkv2 = bf(B, S, 3, H, dh)
    cache = F_ref.rope_cache(S, dh, device=dev)
    ids = torch.randint(0, 10000, (B, S), device=dev)
    emb, demb = torch.randn(10000, d, device=dev), torch.empty(10000, d, device=dev)
    grads = [torch.randn(dff, d, device=dev) for _ in range(8)]
    f32 = torch.randn(T, dff, device=dev)
    for _ in range(reps):
        k.rope_qk_inplace(qkv2, cache)
        e = k.embedding_fwd(ids, emb)
        k.embedding_bwd(ids, e, demb)
        k.clip_grad_norm(grads, 1.0, 1e-6)
        k.cast_bf16(f32)
        k.cast3_bf16(f32, f32.clone(), f32.clone())
        k.residual_add(torch.randn(T, d, device=dev), bf(T, d))
if "comm" in which:
    # the NVLink kernels on ONE rank (they meet themselves at the barriers): memory-side behaviour only, no link traffic
    import ctypes
    import torch.distributed as dist
    from cs336_a2_systems_b200 import _native
    from cs336_a2_systems_b200.parallel import symm
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1"); os.environ.setdefault("MASTER_PORT", "29877")
    dist.init_process_group("nccl", rank=0, world_size=1, device_id=torch.device("cuda", 0))
    n = 64 * 1024 * 1024 // 4
    ga, pa, sa = (symm.SymmArena(n * 4 + 256, torch.device("cuda", 0), None) for _ in range(3))
    ga.view(0, n, torch.float32).normal_(); pa.view(0, n, torch.float32).normal_()
    algos = [1, 2] + ([3] if ga.multicast_ptr else [])
    m_, v_ = torch.zeros(n + 8, device=dev), torch.zeros(n + 8, device=dev)
    hyper = torch.tensor([1e-4, 0.9, 0.95, 1e-8, 0.1, 1.0, 0, 0], device=dev)
    step = torch.ones(1, dtype=torch.int32, device=dev)
    none8 = (ctypes.c_longlong * 8)()
    mcs = [ga.multicast_ptr, pa.multicast_ptr, sa.multicast_ptr]
    mc = k.host_ptr_array(mcs) if all(mcs) else None
    table = torch.tensor([[0, (n * 4 + 31) // 32]], dtype=torch.int64, device=dev)
    st = torch.cuda.current_stream().cuda_stream
    for _ in range(reps):
        for algo in algos:
            k.allreduce_avg(ga, 0, 16384 if algo == 1 else n, torch.float32, 0, 0, algo=algo)
        k.bcast_pull(ga, 0, n * 4, 0)
        if ga.multicast_ptr:
            k.bcast_mc(ga, 0, n * 4, 0)
        for variant in ([0, 3] if mc is not None else [0]):
            for defer in (0, 1):
                _native.check(_native.lib().b200_zero_adamw_bucket(
                    ga.peer_ptrs_c, pa.peer_ptrs_c, sa.peer_ptrs_c, ga.signal_ptrs_c, mc, ctypes.c_longlong(0),
                    ctypes.c_longlong(n), _native.p(m_), _native.p(v_), _native.p(hyper), _native.p(step), ctypes.c_int(0),
                    ctypes.c_int(0), ctypes.c_int(1), ctypes.c_int(0), ctypes.c_int(0), ctypes.c_int(variant),
                    ctypes.c_int(defer), ctypes.c_int(0), none8, none8, ctypes.c_void_p(st)), "zero")
        _native.check(_native.lib().b200_zero_param_gather(pa.peer_ptrs_c, ga.signal_ptrs_c, ctypes.c_ulonglong(0),
                                                           _native.p(table), ctypes.c_int(1), ctypes.c_int(0), ctypes.c_int(1),
                                                           ctypes.c_int(0), ctypes.c_int(0), ctypes.c_void_p(st)), "gather")
        launch = k.AdamWLaunch([pa.view(0, n, torch.float32)], [ga.view(0, n, torch.float32)], [m_[:n]], [v_[:n]],
                               [sa.view(0, n, torch.bfloat16)], step0=1)
        launch.set_hyper(1e-4, 0.9, 0.95, 1e-8, 0.1)
        k.adamw_multi_distribute(launch, pa, sa)
        k.symm_barrier(ga, 0)
    torch.cuda.synchronize()
    k.comm_check()
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
