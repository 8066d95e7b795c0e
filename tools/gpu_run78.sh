#!/bin/bash
# run 78: full GPU suite + smoke + default bench lines on the tree with the async save and the checkpoint API
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_78.log 2>&1; echo rc=$? >> gpurun_out/pytest_78.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_78.log 2>&1; echo rc=$? >> gpurun_out/smoke_78.log
timeout 900 python bench.py > gpurun_out/bench_default_78.log 2>&1; echo rc=$?
timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref_78.log 2>&1; echo rc=$?
