#!/bin/bash
# Builds the reference's own hot-path sources, unmodified and where they lie under
# /root/reference, into oracle/_ref/ (git-ignored, travels to the GPU box):
#   libvulcan_ref.so         -O3 -DNDEBUG, default -fmad=true  (the timing baseline and parity target)
#   libvulcan_ref_nofmad.so  same with -fmad=false             (contraction-free parity target)
#   libvulcan_ref_asshipped.so  -O3 WITHOUT -DNDEBUG: what the reference's own CMake produces
#                            (CMakeLists.txt:36-38 sets no build type): a cudaDeviceSynchronize after
#                            every launch (include/vulcan/device.h:45-52) and live device-side checks;
#                            timed once for the record, never the bar
# Only nvcc on the few source files; the reference's CMake build is not run (needs
# OpenCV / Eigen / HAL / gflags / glog, all absent).  OpenCV is replaced by
# oracle/ref_stub/opencv2/opencv.hpp, the few Eigen types of the tracker by oracle/ref_stub/Eigen/Eigen; `-include cfloat` supplies FLT_MAX for
# src/light_integrator.cu:79-80.
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${VULCAN_REFERENCE:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/src" ]; then echo "build_ref.sh: $REF not present, keeping prebuilt $OUT"; exit 0; fi
mkdir -p "$OUT"
SRCS="volume.cu integrator.cu depth_integrator.cu color_integrator.cu light_integrator.cu tracer.cu tracer.cpp frame.cu frame.cpp image.cu tracker.cpp depth_tracker.cpp depth_tracker.cu"
FILES=""
for s in $SRCS; do FILES="$FILES $REF/src/$s"; done
BASE="-std=c++17 -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -include cfloat -x cu -w \
  -I$HERE/ref_stub -I$REF/include -Xcompiler -fPIC -Xlinker -Bsymbolic -shared -lcublas"
COMMON="$BASE -DNDEBUG"
nvcc $COMMON -o "$OUT/libvulcan_ref.so" $FILES "$HERE/ref_shim.cu" &
nvcc $COMMON -fmad=false -o "$OUT/libvulcan_ref_nofmad.so" $FILES "$HERE/ref_shim.cu" &
nvcc $BASE -o "$OUT/libvulcan_ref_asshipped.so" $FILES "$HERE/ref_shim.cu" &
wait
ls -la "$OUT"
