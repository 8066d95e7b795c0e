set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_70.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_70.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_70.log 2>&1; echo "rc=$?" >> gpurun_out/smoke_70.log
timeout 900 python bench.py > gpurun_out/bench_default_70.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_70.log
timeout 900 python bench.py --impl reference > gpurun_out/bench_ref_70.log 2>&1; echo "rc=$?" >> gpurun_out/bench_ref_70.log
