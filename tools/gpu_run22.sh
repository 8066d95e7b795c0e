set -x
mkdir -p gpurun_out
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_a.log 2>&1
NOVUM_MERGE_BIN_PRIMARY=1 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_bin0.log 2>&1
