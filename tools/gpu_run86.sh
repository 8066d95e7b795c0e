#!/bin/bash
# run 86: ncu --set full of the radix kernels after the row-scan change (scatterPass now scans the digit totals itself; digitRowScan is new)
set -x
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"scatterPass|digitRowScan|tableKernel" -s 30 -c 10 -o gpurun_out/grid_86 python bench.py --workload hashgrid --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_grid_86.log 2>&1; echo rc=$?
