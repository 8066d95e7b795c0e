set -x
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_a.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_b.log 2>&1
