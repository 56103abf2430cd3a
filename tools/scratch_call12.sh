cd /root/repo
timeout 1200 python -m pytest tests -x -q -m gpu -p no:cacheprovider > gpurun_out/c12_pytest.log 2>&1; echo "[pytest] rc=$? $(tail -1 gpurun_out/c12_pytest.log)"
for cfg in "default:" "noswiglu:B200SYS_FUSE_SWIGLU=0"; do
  name=${cfg%%:*}; envs=${cfg#*:}
  env $envs timeout 200 python bench.py --steps 10 --warmup 3 > gpurun_out/c12_bench_$name.log 2> gpurun_out/c12_bench_$name.err
  echo "[$name] $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/c12_bench_$name.log | head -1)"
done
B200SYS_PDL=0 timeout 200 python tools/profile_step.py --out gpurun_out/step_profile_xl_eager_nopdl.json > gpurun_out/c12_profile.log 2>&1
head -16 gpurun_out/c12_profile.log
