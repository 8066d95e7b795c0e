set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/b2_vcm_61.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_61.log
timeout 600 $TR bench.py --gpus 2 --steps 4 --warmup 3 --sharding tiles > gpurun_out/b2_vcm_tiles_61.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_tiles_61.log
timeout 600 $TR bench.py --impl reference --gpus 2 --steps 2 --warmup 1 > gpurun_out/b2_ref_61.log 2>&1; echo "rc=$?" >> gpurun_out/b2_ref_61.log
