set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/b2_vcm_striped.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_striped.log
timeout 900 $TR bench.py --gpus 2 --steps 4 --warmup 3 --sharding tiles > gpurun_out/b2_vcm_tiles.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_tiles.log
timeout 900 python bench.py --steps 4 --warmup 3 --sharding tiles --no-cpu-baseline > gpurun_out/b1_vcm_tiles.log 2>&1; echo "rc=$?" >> gpurun_out/b1_vcm_tiles.log
timeout 900 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload dragon_nee > gpurun_out/b2_dragon.log 2>&1; echo "rc=$?" >> gpurun_out/b2_dragon.log
timeout 900 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload water_sppm --sharding tiles > gpurun_out/b2_sppm_tiles.log 2>&1; echo "rc=$?" >> gpurun_out/b2_sppm_tiles.log
timeout 900 $TR bench.py --gpus 2 --impl reference --steps 1 --warmup 1 --workload cornell > gpurun_out/b2_ref.log 2>&1; echo "rc=$?" >> gpurun_out/b2_ref.log
