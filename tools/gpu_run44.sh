set -x
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/gpus.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_44.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_44.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_44.log 2>&1; echo "rc=$?" >> gpurun_out/smoke_44.log
timeout 900 python bench.py > gpurun_out/bench_default_44.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_44.log
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_44.log 2>&1; echo "rc=$?" >> gpurun_out/bench_ref_44.log
timeout 900 python bench.py --workload dragon_nee --no-cpu-baseline > gpurun_out/bench_dragon_44.log 2>&1; echo "rc=$?" >> gpurun_out/bench_dragon_44.log
tail -3 gpurun_out/pytest_gpu_44.log gpurun_out/smoke_44.log gpurun_out/bench_default_44.log
