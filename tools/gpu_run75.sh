#!/bin/bash
# run 75: A/B of the branch-free accept path in the Lambert merge sweep
set -x
mkdir -p gpurun_out
L=cudapathtracer_b200/libnovum_b200.so
timeout 900 python -m pytest tests/test_gpu_bidir.py -x -q -m gpu > gpurun_out/pytest_75.log 2>&1; echo rc=$? >> gpurun_out/pytest_75.log
for v in flat noflat flat noflat; do
  cp build_alt/lib_$v.so $L
  timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu-baseline >> gpurun_out/bench_vcm_75_$v.log 2>&1; echo rc=$?
done
cp build_alt/lib_flat.so $L
timeout 600 python bench.py --workload water_sppm --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sppm_75_flat.log 2>&1
cp build_alt/lib_noflat.so $L
timeout 600 python bench.py --workload water_sppm --steps 4 --warmup 3 --no-cpu-baseline > gpurun_out/bench_sppm_75_noflat.log 2>&1
cp build_alt/lib_flat.so $L
