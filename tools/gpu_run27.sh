set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511"
timeout 900 $TR bench.py --gpus 2 --steps 4 --warmup 3 --sharding tiles --exchange p2p > gpurun_out/b2_vcm_tiles_p2p.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_tiles_p2p.log
timeout 900 $TR bench.py --gpus 2 --steps 4 --warmup 3 --sharding tiles --exchange nccl > gpurun_out/b2_vcm_tiles_nccl.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_tiles_nccl.log
timeout 900 $TR bench.py --gpus 2 --steps 4 --warmup 3 > gpurun_out/b2_vcm_striped.log 2>&1; echo "rc=$?" >> gpurun_out/b2_vcm_striped.log
timeout 900 $TR bench.py --gpus 2 --steps 3 --warmup 3 --workload water_sppm --sharding tiles > gpurun_out/b2_sppm_tiles.log 2>&1; echo "rc=$?" >> gpurun_out/b2_sppm_tiles.log
