#!/bin/bash
# run 82: full GPU suite on the tree with the chunked multi-threaded OBJ front end + its timing on the box's host cores
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/ -x -q -m gpu > gpurun_out/pytest_82.log 2>&1; echo rc=$? >> gpurun_out/pytest_82.log
nproc > gpurun_out/obj_timing_82.log
./cudapathtracer_b200/novum --make-scene dragon --out /tmp/dragon100 --detail 100 > /dev/null
for r in 1 2 3; do for a in "1 1073741824" "0 0" "8 0" "16 262144"; do python tools/obj_parse_timing.py /tmp/dragon100/scenedata/dragon_vrip.obj $a >> gpurun_out/obj_timing_82.log 2>&1; done; done
timeout 600 python tools/build_timing.py > gpurun_out/build_timing_82.log 2>&1; echo rc=$?
timeout 600 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_82.log 2>&1; echo rc=$? >> gpurun_out/smoke_82.log
