set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_hashgrid.py tests/test_gpu_trace.py -m gpu -q -x > gpurun_out/pytest_59.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_59.log
timeout 900 python bench.py --workload hashgrid --steps 3 --warmup 1 > gpurun_out/bench_grid_59p.log 2>&1; echo "rc=$?" >> gpurun_out/bench_grid_59p.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid_59.csv python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_grid_59.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"scatterPass|hashKernel" -s 24 -c 6 -o gpurun_out/grid_59 python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_grid_59.log 2>&1
