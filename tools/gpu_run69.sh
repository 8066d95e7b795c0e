set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/pytest_69.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_69.log
timeout 900 python bench.py --workload trace_batch --steps 10 --warmup 2 --no-cpu-baseline > gpurun_out/bench_trace_69.log 2>&1
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_69.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_69.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell_69.log 2>&1
timeout 900 python bench.py --workload water_bdpt --no-cpu-baseline > gpurun_out/bench_bdpt_69.log 2>&1
timeout 600 python tools/build_timing.py > gpurun_out/build_timing_69.log 2>&1
