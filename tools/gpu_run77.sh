#!/bin/bash
# run 77: asynchronous frame save (tests + bench numbers)
set -x
mkdir -p gpurun_out
true
timeout 600 python bench.py --workload frame_finish --warmup 3 > gpurun_out/bench_frame_77.log 2>&1; echo rc=$?
