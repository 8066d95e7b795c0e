set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus8.txt
nvidia-smi topo -m > gpurun_out/topo8.txt 2>&1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511"
timeout 600 $TR bench.py --gpus 8 --steps 4 --warmup 3 > gpurun_out/b8_vcm_striped.log 2>&1; echo "rc=$?" >> gpurun_out/b8_vcm_striped.log
timeout 600 $TR bench.py --gpus 8 --steps 4 --warmup 3 --sharding tiles --exchange p2p > gpurun_out/b8_vcm_tiles_p2p.log 2>&1; echo "rc=$?" >> gpurun_out/b8_vcm_tiles_p2p.log
timeout 600 $TR bench.py --gpus 8 --steps 4 --warmup 3 --sharding tiles --exchange nccl > gpurun_out/b8_vcm_tiles_nccl.log 2>&1; echo "rc=$?" >> gpurun_out/b8_vcm_tiles_nccl.log
timeout 600 $TR bench.py --gpus 8 --steps 3 --warmup 3 --workload dragon_nee > gpurun_out/b8_dragon.log 2>&1; echo "rc=$?" >> gpurun_out/b8_dragon.log
TR4="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29512"
timeout 600 $TR4 bench.py --gpus 4 --steps 4 --warmup 3 > gpurun_out/b4_vcm_striped.log 2>&1; echo "rc=$?" >> gpurun_out/b4_vcm_striped.log
timeout 600 $TR4 bench.py --gpus 4 --steps 4 --warmup 3 --sharding tiles > gpurun_out/b4_vcm_tiles_p2p.log 2>&1; echo "rc=$?" >> gpurun_out/b4_vcm_tiles_p2p.log
