#!/bin/bash
# run 85: final tree - full GPU suite, smoke, default bench
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_85.log 2>&1; echo rc=$? >> gpurun_out/pytest_85.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_85.log 2>&1; echo rc=$? >> gpurun_out/smoke_85.log
timeout 900 python bench.py > gpurun_out/bench_default_85.log 2>&1; echo rc=$?
