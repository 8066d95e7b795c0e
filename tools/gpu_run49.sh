set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_frame.py tests/test_gpu_bidir.py -m gpu -q -x > gpurun_out/pytest_49.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_49.log
timeout 600 python bench.py --workload frame_finish --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_frame_49.log 2>&1; echo "rc=$?" >> gpurun_out/bench_frame_49.log
timeout 900 python bench.py --no-cpu-baseline > gpurun_out/bench_default_49.log 2>&1; echo "rc=$?" >> gpurun_out/bench_default_49.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:frameFinish -s 5 -c 2 -o gpurun_out/frame_finish_49 python bench.py --workload frame_finish --steps 3 --warmup 5 --no-cpu-baseline > gpurun_out/ncu_frame_49.log 2>&1
