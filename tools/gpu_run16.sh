set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_a.log 2>&1
NOVUM_MERGE_CTAS=5 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_c5.log 2>&1
NOVUM_MERGE_CTAS=4 timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_c4.log 2>&1
timeout 900 python bench.py --workload water_sppm --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_sppm.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bdMergeLambert -s 15 -c 2 -o gpurun_out/prof_mergeLambert5 $CMD > gpurun_out/ncu_merge.log 2>&1
