set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
timeout 900 python bench.py --workload dragon_nee --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_nee.log 2>&1
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_a.log 2>&1
timeout 900 python bench.py --workload cornell --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_cornell.log 2>&1
timeout 900 python bench.py --workload water_bdpt --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_bdpt.log 2>&1
timeout 900 python bench.py --workload water_sppm --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_sppm.log 2>&1
