set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"scatterPass|digitHistogram|hashKernel|tableKernel|scanApply" -s 60 -c 12 -o gpurun_out/grid_56 python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_grid_56.log 2>&1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_grid_56.csv python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_l_grid_56.log 2>&1
