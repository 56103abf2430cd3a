#!/bin/bash
# The BASELINE.json matrix on an N-GPU box, both arms, one time-limited torchrun per cell (logs: gpurun_out/mx_*.json):
#   buckets   XL bucketed DDP at bucket_size_mb 1 / 10 / 100 / 1000   (reference: ...sharded.py:138-142)
#   perparam  XL per-parameter DDP                                     (reference: ...sharded.py:217-248)
#   sharded27 2.7B DDP + sharded optimizer, step time + peak memory    (reference: ...sharded.py:143-192)
# usage: tools/bench_matrix.sh N "buckets perparam sharded27" "ours reference" [bucket sizes]
set -u
N=${1:-2}; WHAT=${2:-"buckets"}; ARMS=${3:-"ours reference"}; SIZES=${4:-"1 10 100 1000"}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29800
cell() {  # name tmo args...
  local name=$1 tmo=$2; shift 2
  port=$((port + 1))
  timeout -s TERM -k 10 "$tmo" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port $port bench.py --gpus "$N" --steps 10 --warmup 3 --watchdog-s $((tmo - 20)) "$@" \
    > "gpurun_out/mx_${name}_n$N.json" 2> "gpurun_out/mx_${name}_n$N.err"
  echo "[$name n=$N] rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"peak_mem_gib": [0-9.]*\|"device_mem_used_gib_max": [0-9.]*\|"unavailable[^}]*\|"error[^}]*' gpurun_out/mx_${name}_n$N.json | tr '\n' ' ')"
}
for arm in $ARMS; do
  for w in $WHAT; do
    case $w in
      buckets) for mb in $SIZES; do cell ${arm}_bucket${mb} 300 --impl $arm --bucket-mb $mb; done ;;
      perparam) cell ${arm}_perparam 300 --impl $arm --ddp individual ;;
      sharded27) cell ${arm}_27b_sharded 420 --impl $arm --model 2.7b --sharded ;;
    esac
  done
done
