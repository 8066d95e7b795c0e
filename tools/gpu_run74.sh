#!/bin/bash
# run 74: A/B of the binned extend queue (origin cell x direction octant) on the wavefront integrator; also re-validates the radix row scan
set -x
mkdir -p gpurun_out
L=cudapathtracer_b200/libnovum_b200.so
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_hashgrid.py -x -q -m gpu > gpurun_out/pytest_74.log 2>&1; echo rc=$? >> gpurun_out/pytest_74.log
for v in bins nobins; do
  cp build_alt/lib_$v.so $L
  timeout 600 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_74_$v.log 2>&1; echo rc=$?
  timeout 600 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_74_$v.log 2>&1; echo rc=$?
done
cp build_alt/lib_bins.so $L
timeout 600 python bench.py --workload hashgrid --steps 5 --warmup 2 --no-cpu-baseline > gpurun_out/bench_grid_74.log 2>&1; echo rc=$?
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid_74.csv python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_grid_74.log 2>&1
