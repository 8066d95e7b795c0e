set -x
mkdir -p gpurun_out
timeout 900 python bench.py --workload hashgrid --steps 5 --warmup 2 > gpurun_out/bench_grid_55.log 2>&1; echo "rc=$?" >> gpurun_out/bench_grid_55.log
timeout 900 python bench.py --workload hashgrid --impl reference > gpurun_out/bench_grid_ref_55.log 2>&1; echo "rc=$?" >> gpurun_out/bench_grid_ref_55.log
timeout 900 python bench.py --workload hashgrid --grid-full --steps 3 --warmup 1 --no-cpu-baseline > gpurun_out/bench_grid_full_55.log 2>&1; echo "rc=$?" >> gpurun_out/bench_grid_full_55.log
