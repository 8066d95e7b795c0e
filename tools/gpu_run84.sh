#!/bin/bash
# run 84: OBJ text read by several threads into an uninitialised buffer; timing breakdown on the box + scene-loading tests
set -x
mkdir -p gpurun_out
./cudapathtracer_b200/novum --make-scene dragon --out /tmp/dragon100 --detail 100 > /dev/null
nproc > gpurun_out/obj_timing_84.log
for r in 1 2 3; do for a in "1 1073741824" "0 0"; do NOVUM_OBJ_TIMING=1 python tools/obj_parse_timing.py /tmp/dragon100/scenedata/dragon_vrip.obj $a >> gpurun_out/obj_timing_84.log 2>&1; done; done
timeout 600 python tools/build_timing.py > gpurun_out/build_timing_84.log 2>&1; echo rc=$?
timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_trace.py -x -q -m gpu > gpurun_out/pytest_84.log 2>&1; echo rc=$? >> gpurun_out/pytest_84.log
