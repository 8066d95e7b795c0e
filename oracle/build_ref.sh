#!/bin/bash
# Builds, for sm_100a, under oracle/_ref/ only (git-ignored, not gpurun-ignored):
#   novum_ref         the reference's own CUDA code (unmodified, included from /root/reference) + the driver in oracle/ref/
#   novum_ref_count   the same with its two intersection routines wrapped by counting shims (rays per pass for Mrays/s; never timed)
#   novum_ref_nosave  the same with the constant DO_PROGRESSIVERENDER (util.cuh:33) read as false inside deviceCode.cu: the BDPT / VCM
#                     launchers then never stop for the 5-second render.bmp + CSV dump (deviceCode.cu:1893, 3447)
#   novum_dropin      the reference's own main.cu + imageUtil.cu with deviceCode.cu replaced by integration/novum_shim.cu, linked
#                     against cudapathtracer_b200/libnovum_b200.so: the proof of the drop-in boundary
# No-op (exit 0) when /root/reference is absent (GPU box): the prebuilt binaries travel with the snapshot.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REPO="$(dirname "$HERE")"
REF="${NOVUM_REFERENCE_DIR:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -f "$REF/deviceCode.cu" ]; then
  echo "build_ref.sh: $REF not present; keeping prebuilt $OUT/novum_ref if any"
  exit 0
fi
mkdir -p "$OUT"
NVCC="${NVCC:-nvcc}"
# default nvcc flags otherwise (-fmad=true, -prec-div=true): the reference ships no build file
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -w -I$REF"
REFDEPS="$HERE/ref/ref_all.cu $HERE/ref/ref_driver.inl $HERE/ref/ref_scene.inl $HERE/ref/ref_render.inl $REF/deviceCode.cu $REF/main.cu"
LIB="$REPO/cudapathtracer_b200/libnovum_b200.so"
DROPDEPS="$HERE/ref/dropin_all.cu $HERE/ref/ref_scene.inl $HERE/ref/ref_render.inl $REPO/integration/novum_shim.cu $REPO/include/novum_b200.h $REF/main.cu"
stale() { local t="$1"; shift; [ ! -x "$t" ] && return 0; for d in "$@"; do [ "$d" -nt "$t" ] && return 0; done; return 1; }
built=""
if stale "$OUT/novum_ref" $REFDEPS; then $NVCC $FLAGS -o "$OUT/novum_ref" "$HERE/ref/ref_all.cu" -lcudart & built="$built novum_ref"; fi
if stale "$OUT/novum_ref_count" $REFDEPS; then $NVCC $FLAGS -DNOVUM_REF_COUNT_RAYS -o "$OUT/novum_ref_count" "$HERE/ref/ref_all.cu" -lcudart & built="$built novum_ref_count"; fi
if stale "$OUT/novum_ref_nosave" $REFDEPS; then $NVCC $FLAGS -DNOVUM_REF_NOSAVE -o "$OUT/novum_ref_nosave" "$HERE/ref/ref_all.cu" -lcudart & built="$built novum_ref_nosave"; fi
if [ -f "$LIB" ] && stale "$OUT/novum_dropin" $DROPDEPS; then
  $NVCC $FLAGS -I"$REPO/include" -o "$OUT/novum_dropin" "$HERE/ref/dropin_all.cu" -L"$REPO/cudapathtracer_b200" -lnovum_b200 \
        -Xlinker -rpath='$ORIGIN/../../cudapathtracer_b200' -lcudart & built="$built novum_dropin"
fi
wait
[ -x "$OUT/novum_ref" ] && [ -x "$OUT/novum_ref_count" ] && [ -x "$OUT/novum_ref_nosave" ]
[ ! -f "$LIB" ] || [ -x "$OUT/novum_dropin" ]
if [ -n "$built" ]; then echo "build_ref.sh: built$built"; else echo "build_ref.sh: $OUT is up to date"; fi
