set -x
mkdir -p gpurun_out
CMD="python bench.py --workload rainbow_vcm --steps 1 --warmup 3 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/plain_vcm.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bdMerge -s 0 -c 2 -o gpurun_out/prof_merge $CMD > gpurun_out/ncu_merge.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_vcm.csv $CMD > gpurun_out/ncu_list_vcm.log 2>&1
