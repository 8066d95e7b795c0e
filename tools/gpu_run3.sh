set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_bidir.py -m gpu -q -x -s > gpurun_out/pytest_bidir.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_bidir.log
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_bidir.py > gpurun_out/pytest.log 2>&1; echo "rc=$?" >> gpurun_out/pytest.log
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench.log 2>&1; echo "bench rc=$?" >> gpurun_out/bench.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 3 > gpurun_out/bench_ref.log 2>&1; echo "ref rc=$?" >> gpurun_out/bench_ref.log
