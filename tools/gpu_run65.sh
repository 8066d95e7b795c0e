set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_bench_workloads.py -m gpu -q -x > gpurun_out/pytest_65.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_65.log
timeout 900 python bench.py --workload hashgrid --grid-full --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_grid_full_65.log 2>&1; echo "rc=$?" >> gpurun_out/bench_grid_full_65.log
