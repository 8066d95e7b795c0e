set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"bdEyeStep|bdLightStep" -s 78 -c 4 -o gpurun_out/prof_steps3 $CMD > gpurun_out/ncu_steps.log 2>&1
