set -x
mkdir -p gpurun_out
CMD="python bench.py --workload dragon_nee --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_dragon.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:wfShade -s 10 -c 3 -o gpurun_out/prof_shade $CMD > gpurun_out/ncu_shade.log 2>&1
