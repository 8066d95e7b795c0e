set -x
mkdir -p gpurun_out
timeout 1200 python tests/golden/make_golden_scenes.py > gpurun_out/golden_scenes.log 2>&1; echo "rc=$?" >> gpurun_out/golden_scenes.log
cp tests/golden/water_render_*.npz tests/golden/rainbow_render_*.npz gpurun_out/ 2>/dev/null
timeout 1200 python -m pytest tests/test_gpu_bidir.py -m gpu -q -x -s -k delta > gpurun_out/pytest_delta.log 2>&1; echo "rc=$?" >> gpurun_out/pytest_delta.log
