#!/bin/bash
# Compiles csrc/vcm.cu alone and dumps the SASS of one bdMergeLambert instantiation (default MIS=1, MINB=8) to /tmp/merge_sass.txt,
# printing the register / spill line of every instantiation. Offline (no GPU): used to count instructions per sweep trip.
set -e
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
MINB="${1:-8}"
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I"$ROOT/include" -Xptxas -v $EXTRA \
     -c "$ROOT/cudapathtracer_b200/csrc/vcm.cu" -o /tmp/vcm.o 2>&1 | grep -v "^ptxas info" | head -20 || true
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -I"$ROOT/include" -Xptxas -v $EXTRA \
     -c "$ROOT/cudapathtracer_b200/csrc/vcm.cu" -o /tmp/vcm.o 2>&1 | grep -A2 "bdMergeLambertILb1" | grep -E "Used|spill" || true
SYM="_ZN2nv14bdMergeLambertILb1ELi${MINB}EEEvNS_9SceneViewENS_7BdStateENS_8BdParamsENS_10MergeQueueE"
cuobjdump -sass -fun "$SYM" /tmp/vcm.o | grep -E "^\s+/\*[0-9a-f]{4}\*/" | sed -E 's/^\s+\/\*([0-9a-f]+)\*\/\s+/\1 /; s/\s*\/\*.*$//' > /tmp/merge_sass.txt
wc -l /tmp/merge_sass.txt
echo "local ops: $(grep -c 'LDL\|STL' /tmp/merge_sass.txt)"
