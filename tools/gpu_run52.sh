set -x
mkdir -p gpurun_out
timeout 600 python bench.py --workload frame_finish --steps 50 --warmup 5 > gpurun_out/bench_frame_52.log 2>&1; echo "rc=$?" >> gpurun_out/bench_frame_52.log
timeout 600 python bench.py --workload frame_finish > gpurun_out/bench_frame_52b.log 2>&1; echo "rc=$?" >> gpurun_out/bench_frame_52b.log
timeout 900 python bench.py --workload dragon_nee --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/bench_dragon_52.log 2>&1; echo "rc=$?" >> gpurun_out/bench_dragon_52.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/launches_frame_52.csv python bench.py --workload frame_finish --steps 3 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_l_frame_52.log 2>&1
