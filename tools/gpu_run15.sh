set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_g4.log 2>&1
NOVUM_MERGE_G1_DEEP=1 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_g1.log 2>&1
NOVUM_MERGE_CTAS=5 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_g4c5.log 2>&1
timeout 900 python bench.py --workload water_sppm --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_sppm.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_vcm3.csv $CMD > gpurun_out/ncu_list_vcm.log 2>&1
