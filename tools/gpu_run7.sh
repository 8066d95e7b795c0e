set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bidir.py -m gpu -q -s > gpurun_out/pytest_bidir.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bidir.log
for wl in rainbow_vcm water_sppm; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.log 2>&1; echo "rc=$?" >> gpurun_out/bench_$wl.log
done
