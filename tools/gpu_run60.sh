set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu_60.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_gpu_60.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_60.log 2>&1; echo "rc=$?" >> gpurun_out/smoke_60.log
timeout 900 python bench.py > gpurun_out/bench_default_60.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_60.log
timeout 900 python bench.py --workload dragon_nee --no-cpu-baseline > gpurun_out/bench_dragon_60.log 2>&1; echo "rc=$?" >> gpurun_out/bench_dragon_60.log
timeout 900 python bench.py --workload water_sppm --no-cpu-baseline > gpurun_out/bench_sppm_60.log 2>&1; echo "rc=$?" >> gpurun_out/bench_sppm_60.log
timeout 900 python bench.py --workload water_bdpt --no-cpu-baseline > gpurun_out/bench_bdpt_60.log 2>&1; echo "rc=$?" >> gpurun_out/bench_bdpt_60.log
timeout 900 python bench.py --workload frame_finish > gpurun_out/bench_frame_60.log 2>&1; echo "rc=$?" >> gpurun_out/bench_frame_60.log
