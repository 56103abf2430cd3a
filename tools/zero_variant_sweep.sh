#!/bin/bash
# End-to-end sweep of the bucket-fused ZeRO kernel variants through bench.py on N GPUs (default 2), each run in its
# own time-limited torchrun so that a stuck variant cannot take the call with it (see NEXT_STEPS.md §1).
#   usage: tools/zero_variant_sweep.sh [N] [extra bench.py flags...]      e.g. tools/zero_variant_sweep.sh 2 --no-graph
# Output: gpurun_out/zero_sweep_n<N>_v<variant>_<shadow>.json (+ .err)
set -u
N=${1:-2}; shift || true
mkdir -p gpurun_out
cd "$(dirname "$0")/.."
port=29600
for v in 0 1 2 3; do
  for sh in local wire; do
    port=$((port + 1))
    out=gpurun_out/zero_sweep_n${N}_v${v}_${sh}
    # SIGTERM (not KILL) lets torchrun reap its workers; output goes to files, never through a pipe
    B200SYS_ZERO_VARIANT=$v B200SYS_ZERO_SHADOW=$sh timeout -s TERM -k 10 120 \
      python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $port \
      bench.py --gpus "$N" --steps 10 --warmup 4 --optimizer sharded --watchdog-s 100 "$@" > "$out.json" 2> "$out.err"
    rc=$?
    ms=$(grep -o '"ms_per_step": [0-9.]*' "$out.json" | head -1)
    echo "variant=$v shadow=$sh rc=$rc $ms"
  done
done
