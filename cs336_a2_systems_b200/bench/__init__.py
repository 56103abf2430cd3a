"""Benchmark / profiling harness: LM step (``lm``), attention sweep (``attention``), all-reduce (``allreduce``),
GEMM vs cuBLAS (``gemm``), memory snapshots (``memory``), mixed-precision demos (``precision``)."""
