#!/bin/bash
# Second 8-GPU pass: bucket sizes, CTA count of the ZeRO kernel, lazy fp32 replicas, 2.7B DDP + sharded optimizer for
# both arms, reference at its own default bucket size.  Logs: gpurun_out/mx_*.json, gpurun_out/m8_*.log
set -u
N=${1:-8}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29900
cell() { local name=$1 tmo=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  port=$((port + 1))
  env "${envs[@]}" timeout -s TERM -k 10 "$tmo" python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 \
    --master-port $port bench.py --gpus "$N" --steps 10 --warmup 3 --watchdog-s $((tmo - 20)) "$@" \
    > "gpurun_out/mx_${name}_n$N.json" 2> "gpurun_out/mx_${name}_n$N.err"
  echo "[$name n=$N] rc=$? $(grep -o '"ms_per_step": [0-9.]*\|"peak_mem_gib": [0-9.]*\|"device_mem_used_gib_max": [0-9.]*\|"unavailable[^}]*\|"error[^}]*' gpurun_out/mx_${name}_n$N.json | tr '\n' ' ')"
}
port=$((port + 1))
timeout -s TERM -k 10 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $port \
  -m cs336_a2_systems_b200.bench.zero --sizes-mb 100 --blocks 16 32 64 148 --json gpurun_out/zero_kernel_blocks_n$N.json > gpurun_out/m8_zero_blocks.log 2>&1
echo "[zero blocks] rc=$?"
cell ours_default 200 B200SYS_X=1 --
cell ours_blocks64 200 B200SYS_ZERO_BLOCKS=64 --
cell ours_lazy 200 B200SYS_X=1 -- --optimizer sharded --defer-gather lazy
cell ours_bucket1000 200 B200SYS_X=1 -- --bucket-mb 1000
cell ours_27b_sharded 400 B200SYS_X=1 -- --model 2.7b --sharded
cell reference_27b_sharded 500 B200SYS_X=1 -- --impl reference --model 2.7b --sharded
cell reference_bucket1000 400 B200SYS_X=1 -- --impl reference --bucket-mb 1000
