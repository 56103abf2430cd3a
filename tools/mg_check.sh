#!/bin/bash
# Multi-GPU validation pass (run on an N-GPU box): NVLink kernel tests, ZeRO kernel micro-benchmark, end-to-end bench
# with each ZeRO kernel variant x deferred gather.  Every step is time-limited and logs to gpurun_out/mg_*.log.
#   usage: tools/mg_check.sh N [quick]
set -u
N=${1:-2}; MODE=${2:-full}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
port=29700
run() {  # name timeout env... -- cmd...
  local name=$1 tmo=$2; shift 2
  local envs=()
  while [ "$1" != "--" ]; do envs+=("$1"); shift; done
  shift
  env "${envs[@]}" timeout -s TERM -k 10 "$tmo" "$@" > "gpurun_out/mg_$name.log" 2> "gpurun_out/mg_$name.err"
  local rc=$?
  echo "[$name] rc=$rc $(grep -E 'passed|failed|error|"ms_per_step"' gpurun_out/mg_$name.log | tail -3 | cut -c1-400 | tr '\n' ' ')"
  [ $rc -ne 0 ] && tail -5 "gpurun_out/mg_$name.err" | cut -c1-300
}
trun() { port=$((port + 1)); echo python -m torch.distributed.run --nnodes=1 --nproc-per-node "$N" --master-addr 127.0.0.1 --master-port $port; }
nvidia-smi topo -m > gpurun_out/mg_topo.log 2>&1
run pytest_mg 900 B200SYS_X=1 -- python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider
run zero_kernel 300 B200SYS_X=1 -- $(trun) -m cs336_a2_systems_b200.bench.zero --sizes-mb 100 1000 --json gpurun_out/zero_kernel_n$N.json
run allreduce 300 B200SYS_X=1 -- $(trun) -m cs336_a2_systems_b200.bench.allreduce --sizes-mb 0.064 1 10 100 1000 --json gpurun_out/allreduce_n$N.json
IFS=';' read -ra CFGS <<< "${MG_CFGS:-auto auto;0 0;0 1;3 0;3 1;1 0}"   # "<zero variant> <defer>" pairs
for cfg in "${CFGS[@]}"; do
  set -- $cfg
  v=$1; d=$2
  vflag=""; [ "$v" != "auto" ] && vflag="--zero-variant $v"
  run bench_v${v}_d${d} 200 B200SYS_X=1 -- $(trun) bench.py --gpus $N --steps 10 --warmup 3 --optimizer sharded --defer-gather $d $vflag --watchdog-s 170
  [ "$MODE" = quick ] && break
done
run bench_replicated 200 B200SYS_X=1 -- $(trun) bench.py --gpus $N --steps 10 --warmup 3 --optimizer replicated --watchdog-s 170
