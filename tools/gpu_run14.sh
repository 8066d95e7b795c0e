set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
for k in 8 5 4 12; do
NOVUM_MERGE_CTAS=$k timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_c$k.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_c$k.log
done
timeout 900 python bench.py --workload water_sppm --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_water_sppm.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bdMergeLambert -s 15 -c 2 -o gpurun_out/prof_mergeLambert4 $CMD > gpurun_out/ncu_merge.log 2>&1
