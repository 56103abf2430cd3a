#!/bin/bash
# One-GPU validation pass over everything that was written without hardware access (see NEXT_STEPS.md):
# each step runs in its own time-limited process and logs to gpurun_out/exp_*.log, so a hang costs one timeout.
#   usage (on the GPU box):  tools/validate_experimental.sh          # ~4-6 minutes
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {  # name timeout env... -- cmd...
  local name=$1 tmo=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  env "${envs[@]}" timeout -s TERM -k 5 "$tmo" "$@" > "gpurun_out/exp_$name.log" 2>&1
  local rc=$?
  echo "[$name] rc=$rc $(grep -E 'PARITY_|causal: fwd|passed|failed|ms_per_step' gpurun_out/exp_$name.log | tail -4 | tr '\n' ' ' | cut -c1-300)"
}
# 1. FlashAttention variants: parity against the fp32 oracle + timings (default kernels first, as the reference point)
run fa_default 90 B200SYS_FA_FWD_WG2=0 -- python tools/fa_variants_check.py
for v in 1 2 3; do run fa_fwd_wg2_$v 90 B200SYS_FA_FWD_WG2=$v -- python tools/fa_variants_check.py; done
for v in 2 4 2r 4r; do run fa_bwd_wg_$v 90 B200SYS_FA_BWD_WG=$v -- python tools/fa_variants_check.py; done
# 2. gated tests (attention variants through pytest, LR schedule under CUDA graphs)
run gated_tests 400 B200SYS_TEST_EXPERIMENTAL=1 -- python -m pytest tests/test_attention.py tests/test_model_gpu.py -q -m gpu -p no:cacheprovider
# 3. end to end: does the LM step get faster?  (baseline, then each switch)
run bench_base 120 B200SYS_PDL=1 -- python bench.py --steps 10 --warmup 3
run bench_fa_fwd1 120 B200SYS_FA_FWD_WG2=1 -- python bench.py --steps 10 --warmup 3
run bench_fa_bwd2 120 B200SYS_FA_FWD_WG2=1 B200SYS_FA_BWD_WG=2 -- python bench.py --steps 10 --warmup 3
run bench_fa_bwd2r 120 B200SYS_FA_FWD_WG2=1 B200SYS_FA_BWD_WG=2r -- python bench.py --steps 10 --warmup 3
run bench_w13 120 B200SYS_FUSE_W13_WGRAD=1 -- python bench.py --steps 10 --warmup 3
