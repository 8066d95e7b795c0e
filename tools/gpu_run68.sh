set -x
mkdir -p gpurun_out
timeout 900 python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default_68.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_68.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/launches_vcm_68.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_vcm_68.log 2>&1
timeout 900 python bench.py --workload dragon_nee --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_68.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -s 100 -c 1500 --csv --log-file gpurun_out/launches_nee_68.csv python bench.py --workload dragon_nee --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_l_nee_68.log 2>&1
