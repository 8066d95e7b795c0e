#!/bin/bash
# run 80: two-GPU sanity of the final tree (striped + row-sharded VCM, dragon NEE)
set -x
mkdir -p gpurun_out
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $T bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_2gpu_80.log 2>&1; echo rc=$?
timeout 600 $T bench.py --gpus 2 --steps 3 --warmup 3 --no-cpu-baseline --sharding tiles > gpurun_out/bench_2gpu_tiles_80.log 2>&1; echo rc=$?
timeout 600 $T bench.py --gpus 2 --steps 2 --warmup 3 --no-cpu-baseline --workload dragon_nee > gpurun_out/bench_2gpu_dragon_80.log 2>&1; echo rc=$?
