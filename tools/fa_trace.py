"""Phase timing of the FA2 forward softmax loop from in-kernel clock64() stamps (debug hook)."""
import ctypes, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from cs336_a2_systems_b200 import _native
from cs336_a2_systems_b200.ops import kernels as k
lib = _native.lib()
buf = torch.zeros(256, dtype=torch.int64, device="cuda")
for (BH, N, causal) in [(16, 2048, False), (100, 512, True)]:
    q, kk, v = (torch.randn(1, BH, N, 64, device="cuda").bfloat16() for _ in range(3))
    k.fa2_fwd(q, kk, v, causal)
    lib.b200_fa2_fwd_set_trace(ctypes.c_void_p(buf.data_ptr()))
    buf.zero_()
    k.fa2_fwd(q, kk, v, causal)
    torch.cuda.synchronize()
    lib.b200_fa2_fwd_set_trace(ctypes.c_void_p(0))
    t = buf.cpu().tolist()
    t = [x for x in t if x]
    d = [b - a for a, b in zip(t, t[1:])]
    names = ["wait S", "pass1 (max)", "pass2 (exp, P->smem)", "wait PV", "O accumulate"]
    print(f"shape BH={BH} N={N} causal={causal}: {len(t)} stamps; first wait {d[0]} cycles")
    its = (len(d)) // 5
    for j in range(min(its, 6)):
        seg = d[5 * j:5 * j + 5]
        print("  iter", j, {n: c for n, c in zip(names, seg)}, "total", sum(seg))
