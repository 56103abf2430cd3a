#!/bin/bash
# One-GPU evidence pass: full GPU test-suite, per-kernel step profile, ncu --set full over every shipped kernel
# (summarised on the box), FlashAttention sweeps for both arms, the batch-8 attention protocol, the 2.7B memory sweep,
# small-LM bench for both arms.  Every step time-limited; logs and JSON under gpurun_out/ev_*.
set -u
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() { local name=$1 tmo=$2; shift 2
  timeout -s TERM -k 10 "$tmo" "$@" > "gpurun_out/ev_$name.log" 2> "gpurun_out/ev_$name.err"; local rc=$?
  echo "[$name] rc=$rc $(grep -E 'passed|failed|"ms_per_step"' gpurun_out/ev_$name.log | tail -2 | cut -c1-300 | tr '\n' ' ')"
  [ $rc -ne 0 ] && tail -12 "gpurun_out/ev_$name.log" | cut -c1-300 && tail -4 "gpurun_out/ev_$name.err" | cut -c1-300
  return $rc; }
STEPS=${EV_STEPS:-"pytest profile ncu fa_ours fa_ref attn8 memory small bench"}
for st in $STEPS; do case $st in
  pytest) run pytest 1500 python -m pytest tests -x -q -m gpu -p no:cacheprovider ;;
  profile) run profile_eager 300 python tools/profile_step.py --out gpurun_out/step_profile_xl_eager.json ;;
  profile0) B200SYS_PDL=0 run profile_eager_nopdl 300 python tools/profile_step.py --out gpurun_out/step_profile_xl_eager_nopdl.json ;;
  ncu) NCU_REPS=1 run ncu_plain 300 python tools/ncu_targets.py &&
       NCU_REPS=1 run ncu 1200 ncu --set full --clock-control none -o gpurun_out/prof_r2 -f python tools/ncu_targets.py
       python tools/ncu_summary.py gpurun_out/prof_r2.ncu-rep > gpurun_out/ncu_r2_summary.md 2> gpurun_out/ev_ncu_summary.err
       ncu -i gpurun_out/prof_r2.ncu-rep --page raw --csv > gpurun_out/ncu_r2_raw.csv 2>/dev/null
       ls -la gpurun_out/prof_r2.ncu-rep; [ $(stat -c %s gpurun_out/prof_r2.ncu-rep 2>/dev/null || echo 0) -gt 40000000 ] && rm -f gpurun_out/prof_r2.ncu-rep ;;
  fa_ours) run fa_ours 600 python -m cs336_a2_systems_b200.bench.attention --json gpurun_out/fa_sweep_ours.json --md gpurun_out/fa_sweep_ours.md ;;
  fa_ref) run fa_quick_both 600 python -m cs336_a2_systems_b200.bench.attention --quick --with-reference --json gpurun_out/fa_quick_both.json --md gpurun_out/fa_quick_both.md
          run fa_ref 900 python -m cs336_a2_systems_b200.bench.attention --with-reference --iters 3 --warmup 1 --json gpurun_out/fa_sweep_both.json --md gpurun_out/fa_sweep_both.md ;;
  attn8) run attn8 600 python -m cs336_a2_systems_b200.bench.attention --protocol batch8 --iters 5 --json gpurun_out/attention_batch8.json --md gpurun_out/attention_batch8.md ;;
  memory) run memory 600 python -m cs336_a2_systems_b200.bench.memory --size 2.7b --ctx 128 256 512 --json gpurun_out/memory_2p7b.json ;;
  small) run small_ours 200 python bench.py --model small --steps 20 --warmup 5
         run small_ref 300 python bench.py --model small --steps 20 --warmup 5 --impl reference ;;
  bench) run bench 200 python bench.py --steps 10 --warmup 5 ;;
esac; done
