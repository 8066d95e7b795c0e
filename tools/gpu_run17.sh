set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
for k in 8 6 5 7 10; do
NOVUM_MERGE_CTAS=$k timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_c$k.log 2>&1
done
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bdEyeStep -s 48 -c 2 -o gpurun_out/prof_eyeStep $CMD > gpurun_out/ncu_eye.log 2>&1
