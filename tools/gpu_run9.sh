set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
for wl in dragon_nee rainbow_vcm water_bdpt water_sppm; do
  timeout 900 python bench.py --workload $wl --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_$wl.log 2>&1; echo "rc=$?" >> gpurun_out/bench_$wl.log
done
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wfExtend -s 10 -c 3 -o gpurun_out/prof_extend2 $CMD > gpurun_out/ncu_full.log 2>&1
