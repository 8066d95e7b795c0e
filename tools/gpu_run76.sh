#!/bin/bash
# run 76: key bits split evenly over the radix passes (28 bits = 4 x 7)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hashgrid.py tests/test_gpu_bench_workloads.py tests/test_gpu_trace.py -x -q -m gpu > gpurun_out/pytest_76.log 2>&1; echo rc=$? >> gpurun_out/pytest_76.log
timeout 600 python bench.py --workload hashgrid --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/bench_grid_76.log 2>&1; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid_76.csv python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_grid_76.log 2>&1
