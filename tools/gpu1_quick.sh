#!/bin/bash
# Short one-GPU pass for kernel iteration: selected tests, GEMM comparison, bench with and without the optimizer epilogue.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local name=$1 tmo=$2; shift 2
  timeout -s TERM -k 10 "$tmo" "$@" > "gpurun_out/q_$name.log" 2> "gpurun_out/q_$name.err"; local rc=$?
  echo "[$name] rc=$rc $(grep -E 'passed|failed|"ms_per_step"' gpurun_out/q_$name.log | tail -3 | cut -c1-330 | tr '\n' ' ')"
  [ $rc -ne 0 ] && tail -25 "gpurun_out/q_$name.log" | cut -c1-300 && tail -5 "gpurun_out/q_$name.err" | cut -c1-300; }
run tests 600 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py -x -q -m gpu -p no:cacheprovider -k "${QUICK_K:-adamw or softmax or materialized or optimizer_in_backward or load_state or rmsnorm or wgrad or native_lm or swiglu or training}"
run gemm 400 python -m cs336_a2_systems_b200.bench.gemm --model xl --json gpurun_out/gemm_vs_cublas_xl.json
run bench 200 python bench.py --steps 10 --warmup 3
B200SYS_FUSE_SWIGLU=0 run bench_noswiglu 200 python bench.py --steps 10 --warmup 3
grep -E "wgrad" gpurun_out/q_gemm.log | cut -c1-200
