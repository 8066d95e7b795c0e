#!/bin/bash
# Builds the reference's own CUDA code (unmodified, included from /root/reference) + the driver in oracle/ref/
# into oracle/_ref/novum_ref for sm_100a. Output only under oracle/_ref/ (git-ignored, not gpurun-ignored).
# No-op (exit 0) when /root/reference is absent (GPU box): the prebuilt binary travels with the snapshot.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${NOVUM_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/deviceCode.cu" ]; then
  echo "build_ref.sh: $REF not present; keeping prebuilt $OUT/novum_ref if any"
  exit 0
fi
mkdir -p "$OUT"
if [ "$OUT/novum_ref" -nt "$HERE/ref/ref_all.cu" ] && [ "$OUT/novum_ref" -nt "$HERE/ref/ref_driver.inl" ] && [ "$OUT/novum_ref" -nt "$REF/deviceCode.cu" ]; then
  echo "build_ref.sh: $OUT/novum_ref is up to date"; exit 0
fi
NVCC="${NVCC:-nvcc}"
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -w -I$REF"
# default nvcc flags otherwise (-fmad=true, -prec-div=true): the reference ships no build file
$NVCC $FLAGS -o "$OUT/novum_ref" "$HERE/ref/ref_all.cu" -lcudart &
$NVCC $FLAGS -DNOVUM_REF_COUNT_RAYS -o "$OUT/novum_ref_count" "$HERE/ref/ref_all.cu" -lcudart &
wait
[ -x "$OUT/novum_ref" ] && [ -x "$OUT/novum_ref_count" ]
echo "build_ref.sh: built $OUT/novum_ref"
