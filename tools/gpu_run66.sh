set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_66.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_66.log
timeout 900 python bench.py --no-cpu-baseline --steps 2 > gpurun_out/bench_default_66.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_66.log
