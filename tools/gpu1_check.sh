#!/bin/bash
# One-GPU validation pass: the whole GPU test-suite, the FlashAttention parity/timing tool, the GEMM comparison and a
# short bench.py run, each in its own time-limited process (logs: gpurun_out/g1_*.log).
#   usage: tools/gpu1_check.sh [skip-tests]
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # name timeout cmd...
  local name=$1 tmo=$2; shift 2
  timeout -s TERM -k 10 "$tmo" "$@" > "gpurun_out/g1_$name.log" 2> "gpurun_out/g1_$name.err"
  local rc=$?
  echo "[$name] rc=$rc $(grep -E 'passed|failed|PARITY_|"ms_per_step"|causal: fwd' gpurun_out/g1_$name.log | tail -6 | cut -c1-330 | tr '\n' ' ')"
  [ $rc -ne 0 ] && tail -15 "gpurun_out/g1_$name.log" | cut -c1-400 && tail -5 "gpurun_out/g1_$name.err" | cut -c1-300
}
[ "${1:-}" != "skip-tests" ] && run pytest 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider
run fa_check 200 python tools/fa_check.py --json gpurun_out/fa_check.json
run bench 200 python bench.py --steps 10 --warmup 3
run smoke 200 python __graft_entry__.py smoke
run gemm 400 python -m cs336_a2_systems_b200.bench.gemm --model xl --json gpurun_out/gemm_vs_cublas_xl.json
run bench_optbwd 200 python bench.py --steps 10 --warmup 3 --opt-in-bwd
