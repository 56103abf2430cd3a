"""GEMM micro-benchmark: our tcgen05 kernel vs cuBLAS (torch.matmul) on the LM's shapes.

CUDA-event timing, L2 flushed between iterations, reports TFLOP/s and the fraction of the measured
cuBLAS bf16 peak (MEASURED_PEAKS.json).  ``python -m cs336_a2_systems_b200.bench.gemm --model xl``
"""
from __future__ import annotations

import argparse
import json
import os

import torch

from ..models import get_model_config
from ..ops import kernels
from ..utils.timing import time_cuda


def lm_shapes(model: str, tokens: int, vocab: int = 10000):
    c = get_model_config(model)
    d, f = c["d_model"], c["d_ff"]
    layers = [("qkv", 3 * d, d), ("out", d, d), ("w13", 2 * f, d), ("w2", d, f), ("head", vocab, d)]
    out = []
    for name, n, k in layers:
        out.append((f"{name}.fwd", "nt", tokens, n, k))       # y = x W^T
        out.append((f"{name}.dgrad", "nn", tokens, k, n))     # dx = dy W      (M=T, N=d_in, K=d_out)
        out.append((f"{name}.wgrad", "tn", n, k, tokens))     # dW = dy^T x    (M=d_out, N=d_in, K=T)
    return out


def peak_tflops() -> float:
    p = os.path.join(os.path.dirname(__file__), "..", "..", "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(p))["bf16_tflops"])
    except Exception:
        return 1590.0


def bench_one(kind: str, M: int, N: int, K: int, iters: int = 10, force_bn: int = 0, pair: bool = False):
    dev = "cuda"
    if kind == "nt":
        a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(N, K, device=dev).bfloat16()
        ours = lambda: kernels.gemm(a, b, force_bn=force_bn, pair=pair)
        ref = lambda: a @ b.t()
    elif kind == "nn":
        a = torch.randn(M, K, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16()
        ours = lambda: kernels.gemm(a, b, b_mn=True, force_bn=force_bn, pair=pair)
        ref = lambda: a @ b
    else:
        a = torch.randn(K, M, device=dev).bfloat16(); b = torch.randn(K, N, device=dev).bfloat16()
        out = torch.empty(M, N, device=dev)
        ours = lambda: kernels.gemm(a, b, a_mn=True, b_mn=True, epilogue=kernels.EPI_F32, out=out, force_bn=force_bn, pair=pair)
        ref = lambda: a.t() @ b
    # 20 back-to-back launches per timed region: a single short kernel after an idle gap runs at idle clocks
    rep = 20
    t_ours = time_cuda(lambda: [ours() for _ in range(rep)], warmup=2, iters=iters, flush=False)["median_ms"] / rep
    t_ref = time_cuda(lambda: [ref() for _ in range(rep)], warmup=2, iters=iters, flush=False)["median_ms"] / rep
    fl = 2.0 * M * N * K
    return t_ours, t_ref, fl / t_ours / 1e9, fl / t_ref / 1e9


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="xl")
    ap.add_argument("--tokens", type=int, default=2048)
    ap.add_argument("--bn", type=int, default=0)
    ap.add_argument("--pair", action="store_true", help="CTA-pair (cta_group::2) kernel")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    peak = peak_tflops()
    rows = []
    print(f"{'gemm':14s} {'kind':4s} {'M':>6s} {'N':>6s} {'K':>6s} {'ours ms':>9s} {'cublas ms':>9s} {'ours TF':>8s} {'cublas TF':>9s} {'of peak':>7s}")
    for name, kind, M, N, K in lm_shapes(args.model, args.tokens):
        to, tr, fo, fr = bench_one(kind, M, N, K, force_bn=args.bn, pair=args.pair)
        rows.append(dict(name=name, kind=kind, M=M, N=N, K=K, ours_ms=to, cublas_ms=tr, ours_tflops=fo,
                         cublas_tflops=fr, frac_of_measured_peak=fo / peak))
        print(f"{name:14s} {kind:4s} {M:6d} {N:6d} {K:6d} {to:9.4f} {tr:9.4f} {fo:8.1f} {fr:9.1f} {fo / peak:7.2f}")
    if args.json:
        json.dump({"peak_tflops_measured": peak, "rows": rows}, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
