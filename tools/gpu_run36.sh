set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_bidir.py -m gpu -q -x -k full_size > gpurun_out/pytest_full.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_full.log
