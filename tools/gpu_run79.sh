#!/bin/bash
# run 79: render tests incl. the checkpointed command-line run
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_render.py -x -q -m gpu > gpurun_out/pytest_79.log 2>&1; echo rc=$? >> gpurun_out/pytest_79.log
