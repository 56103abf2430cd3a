#!/bin/bash
# Runs groups of GPU tests in separate, individually time-limited processes so that one hung kernel
# cannot take the whole call (or the box) with it.  Logs land in gpurun_out/.
# usage: tools/gpu_check.sh [group ...]   (default: all groups)
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
run() {  # name timeout cmd...
  local name=$1 tmo=$2; shift 2
  echo "=== $name"
  timeout -s KILL "$tmo" "$@" > "gpurun_out/$name.log" 2>&1
  local rc=$?
  echo "rc=$rc $(tail -n 3 gpurun_out/$name.log | tr '\n' ' ' | cut -c1-400)"
}
groups=${@:-"info elem gemm fa model smoke"}
for g in $groups; do
  case $g in
    info) run info 120 bash -c 'nvidia-smi -L; nvidia-smi --query-gpu=clocks.sm,clocks.max.sm,power.draw --format=csv; python -c "import torch;print(torch.cuda.get_device_name(0), torch.cuda.get_device_capability(0))"' ;;
    elem) run elem 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "not gemm" -p no:cacheprovider ;;
    gemm) run gemm_nt 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm_nt" -p no:cacheprovider
          run gemm_other 300 python -m pytest tests/test_kernels_gpu.py -q -m gpu -k "gemm and not gemm_nt" -p no:cacheprovider ;;
    fa) run fa_fwd 300 python -m pytest tests/test_attention.py -q -m gpu -k "forward" -p no:cacheprovider
        run fa_bwd 300 python -m pytest tests/test_attention.py -q -m gpu -k "backward" -p no:cacheprovider
        run fa_shapes 400 python -m pytest tests/test_attention.py -q -m gpu -k "shapes" -p no:cacheprovider ;;
    model) run model 400 python -m pytest tests/test_model_gpu.py -q -m gpu -p no:cacheprovider ;;
    smoke) run smoke 300 python __graft_entry__.py smoke ;;
    all) run pytest_gpu 1200 python -m pytest tests -x -q -m gpu -p no:cacheprovider ;;
    bench) run bench 900 python bench.py --steps 5 --warmup 3 ;;
    *) echo "unknown group $g" ;;
  esac
done
