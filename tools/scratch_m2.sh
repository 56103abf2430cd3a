cd "$(dirname "$0")/.."
mkdir -p gpurun_out
N=${1:-2}
timeout 300 python -m pytest tests/test_kernels_gpu.py tests/test_model_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu -p no:cacheprovider -k "rmsnorm or native_lm or training or no_alloc or oracle" > gpurun_out/m2_pytest_1gpu.log 2>&1; echo "[pytest norm] rc=$? $(tail -1 gpurun_out/m2_pytest_1gpu.log)"
timeout 900 python -m pytest tests/test_multigpu.py -x -q -p no:cacheprovider > gpurun_out/pytest_multigpu_n$N.log 2>&1; echo "[pytest multigpu] rc=$? $(tail -1 gpurun_out/pytest_multigpu_n$N.log)"
bash tools/bench_matrix.sh $N "buckets perparam" "ours reference" "1 10 100 1000"
