"""FlashAttention-2 benchmark sweep: forward, backward and forward+backward latency.

Counterpart of the reference's attention benchmarks (``benchmark_attention.py:25-201``,
``flashattentioncode.py:15-299``): batch 1, causal, seq 128…65536 × head dim 16…128 × {bf16, fp32},
plus the course leaderboard shape (16 heads × 16384 × 64, bf16, causal).  Implementations:

* ``sm100``     — this package's tcgen05 kernels (``FlashAttentionSM100``)
* ``reference`` — the unmodified reference ``FlashAttentionTriton`` (Triton forward + its backward) from
                  ``baseline/_ref`` when installed (``--with-reference``)
* ``eager``     — materialised PyTorch attention (the reference LM's ``scaled_dot_product_attention``)
* ``materialized`` — the same O(N²) algorithm on this library's kernels (tcgen05 GEMMs + one fused
                  scale/mask/softmax row kernel): the counterpart of the reference's ``torch.compile``d attention arm

``--protocol batch8`` is the reference's second attention benchmark (``benchmark_attention.py:113-201``): batch 8,
seq 256…16384 × d 16…128, forward / backward time and peak memory of the eager, materialized and flash arms.
Every row carries the clocks sampled while it ran.

CUDA-event timing after warm-up, L2 flushed between iterations; OOM is caught and recorded as null
(the reference's backward materialises N×N matrices).  Output: JSON lines + a Markdown table.

    python -m cs336_a2_systems_b200.bench.attention --quick --with-reference --json out.json
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import torch

from ..ops.attention import FlashAttentionSM100
from ..ops.functional import sdpa_ref
from ..utils.clocks import ClockSampler
from ..utils.timing import time_cuda


def _flops(B, H, N, D, causal):
    f = 4.0 * B * H * N * N * D * (0.5 if causal else 1.0)
    return f, 2.5 * f  # fwd, bwd


def _bench_impl(fn_fwd, q, k, v, iters, warmup):
    """Returns (fwd_ms, bwd_ms, fwdbwd_ms) or None triples on OOM."""
    out = {}
    try:
        o = fn_fwd(q, k, v)
        do = torch.randn_like(o)
        out["fwd_ms"] = time_cuda(lambda: fn_fwd(q, k, v), warmup, iters)["median_ms"]

        def bwd():
            q.grad = k.grad = v.grad = None
            o.backward(do, retain_graph=True)

        out["bwd_ms"] = time_cuda(bwd, warmup, iters)["median_ms"]

        def fb():
            q.grad = k.grad = v.grad = None
            fn_fwd(q, k, v).backward(do)

        out["fwdbwd_ms"] = time_cuda(fb, warmup, iters)["median_ms"]
    except torch.OutOfMemoryError:
        out.setdefault("fwd_ms", None)
        out.setdefault("bwd_ms", None)
        out.setdefault("fwdbwd_ms", None)
        out["oom"] = True
        torch.cuda.empty_cache()
    except Exception as e:  # e.g. the reference backward rejects bf16 inputs (fp32 O against bf16 Q/K/V)
        out.setdefault("fwd_ms", None)
        out.setdefault("bwd_ms", None)
        out.setdefault("fwdbwd_ms", None)
        out["error"] = f"{type(e).__name__}: {str(e).splitlines()[-1][:160]}"
    return out


def run(shapes, dtypes, impls, iters, warmup, ref=None):
    rows = []
    dev = "cuda"
    for (BH, N, D) in shapes:
        for dt in dtypes:
            for impl in impls:
                torch.manual_seed(0)
                q, k, v = (torch.randn(BH, N, D, device=dev, dtype=dt, requires_grad=True) for _ in range(3))
                if impl == "sm100":
                    fn = lambda a, b, c: FlashAttentionSM100.apply(a, b, c, True)
                elif impl == "reference":
                    fn = lambda a, b, c: ref.FlashAttentionTriton.apply(a, b, c, True)
                elif impl == "materialized":
                    if N > 16384:
                        continue
                    from ..ops import native
                    fn = lambda a, b, c: native.materialized_attention(a, b, c, True)
                else:
                    mask = torch.ones(N, N, device=dev, dtype=torch.bool).tril() if N <= 16384 else None
                    if mask is None:
                        continue
                    fn = lambda a, b, c: sdpa_ref(a, b, c, mask)
                n_it = max(2, min(iters, int(2e9 / (N * N))))  # keep the eager/reference sweeps bounded
                torch.cuda.reset_peak_memory_stats()
                with ClockSampler(0, period_ms=50) as clk:
                    r = _bench_impl(fn, q, k, v, n_it, warmup)
                peak_mib = round(torch.cuda.max_memory_allocated() / 2 ** 20, 1)
                clocks = clk.summary()
                if not clocks.get("samples") and r.get("fwd_ms"):
                    # the timed region was shorter than one sampling period: record the clocks of the same kernel run
                    # back to back for 0.3 s (a number without a clock record counts as unmeasured)
                    with ClockSampler(0, period_ms=50) as clk2, torch.no_grad():
                        t_end = time.perf_counter() + 0.3
                        while time.perf_counter() < t_end:
                            for _ in range(20):
                                fn(q, k, v)
                            torch.cuda.synchronize()
                    clocks = dict(clk2.summary(), note="sampled over a 0.3 s back-to-back re-run of the forward")
                ff, fb = _flops(1, BH, N, D, True)
                row = dict(impl=impl, heads=BH, seq=N, d=D, dtype=str(dt).split(".")[-1], **r,
                           peak_mem_mib=peak_mib, clocks=clocks)
                if r.get("fwd_ms"):
                    row["fwd_tflops"] = ff / r["fwd_ms"] / 1e9
                if r.get("bwd_ms"):
                    row["bwd_tflops"] = fb / r["bwd_ms"] / 1e9
                rows.append(row)
                print(json.dumps(row), flush=True)
                del q, k, v
                torch.cuda.empty_cache()
    return rows


def markdown(rows):
    out = ["| impl | heads | seq | d | dtype | fwd ms | bwd ms | fwd+bwd ms | fwd TF/s | bwd TF/s | peak MiB | SM MHz | throttle |",
           "|---|---|---|---|---|---|---|---|---|---|---|---|---|"]
    f = lambda x: "n/a" if x is None else (f"{x:.3f}" if isinstance(x, float) else str(x))
    for r in rows:
        out.append(f"| {r['impl']} | {r['heads']} | {r['seq']} | {r['d']} | {r['dtype']} | {f(r.get('fwd_ms'))} | "
                   f"{f(r.get('bwd_ms'))} | {f(r.get('fwdbwd_ms'))} | {f(r.get('fwd_tflops', '-'))} | {f(r.get('bwd_tflops', '-'))} | "
                   f"{f(r.get('peak_mem_mib'))} | {f((r.get('clocks') or {}).get('sm_mhz'))} | "
                   f"{','.join((r.get('clocks') or {}).get('reasons') or []) or '-'} |")
    return "\n".join(out)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--quick", action="store_true", help="a handful of shapes instead of the full sweep")
    ap.add_argument("--with-reference", action="store_true")
    ap.add_argument("--with-eager", action="store_true")
    ap.add_argument("--with-materialized", action="store_true")
    ap.add_argument("--protocol", choices=["sweep", "batch8"], default="sweep")
    ap.add_argument("--dtypes", nargs="+", default=["bfloat16", "float32"])
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--json", default=None)
    ap.add_argument("--md", default=None)
    args = ap.parse_args()
    if args.protocol == "batch8":
        seqs = (256, 4096) if args.quick else (256, 1024, 4096, 8192, 16384)
        shapes = [(8, n, d) for n in seqs for d in ((64,) if args.quick else (16, 32, 64, 128))]
        dtypes = [getattr(torch, d) for d in args.dtypes]
    elif args.quick:
        shapes = [(16, 16384, 64), (1, 4096, 64), (1, 16384, 128), (100, 512, 64)]
        dtypes = [torch.bfloat16, torch.float32]
    else:
        shapes = [(1, n, d) for n in (128, 256, 1024, 4096, 8192, 16384, 32768, 65536) for d in (16, 32, 64, 128)]
        shapes.append((16, 16384, 64))
        dtypes = [torch.bfloat16, torch.float32]
    impls = ["sm100"]
    ref = None
    if args.with_reference:
        sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "..", "baseline"))
        import ref_loader
        ok, why = ref_loader.available()
        if ok:
            ref = ref_loader.load()
            impls.append("reference")
        else:
            print(json.dumps({"reference_unavailable": why}))
    if args.with_eager or args.protocol == "batch8":
        impls.append("eager")
    if args.with_materialized or args.protocol == "batch8":
        impls.append("materialized")
    rows = run(shapes, dtypes, impls, args.iters, args.warmup, ref)
    if args.json:
        json.dump(rows, open(args.json, "w"), indent=1)
    md = markdown(rows)
    if args.md:
        open(args.md, "w").write(md + "\n")
    print(md)


if __name__ == "__main__":
    main()
