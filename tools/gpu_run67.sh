set -x
mkdir -p gpurun_out
timeout 600 python tools/build_timing.py > gpurun_out/build_timing_67.log 2>&1
CUDA_MODULE_LOADING=EAGER timeout 600 python tools/build_timing.py > gpurun_out/build_timing_eager_67.log 2>&1
