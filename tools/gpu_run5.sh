set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
for wl in rainbow_vcm water_sppm; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 > gpurun_out/bench_$wl.log 2>&1; echo "rc=$?" >> gpurun_out/bench_$wl.log
done
