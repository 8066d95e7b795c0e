#!/bin/bash
# run 73: radix sort with the per-digit row scan (one scan kernel per pass instead of three)
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hashgrid.py tests/test_gpu_bench_workloads.py tests/test_gpu_bidir.py -x -q -m gpu > gpurun_out/pytest_73.log 2>&1; echo rc=$? >> gpurun_out/pytest_73.log
timeout 600 python bench.py --workload hashgrid --steps 5 --warmup 2 > gpurun_out/bench_grid_73.log 2>&1; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid_73.csv python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_grid_73.log 2>&1
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/bench_vcm_73.log 2>&1; echo rc=$?
