set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bdEyeStep -s 48 -c 2 -o gpurun_out/prof_eyeStep3 $CMD > gpurun_out/ncu_eye.log 2>&1
