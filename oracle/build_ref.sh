#!/bin/bash
# Builds oracle/_ref/ref_driver: the REFERENCE's own CUDA translation units, compiled from where they lie
# under /root/reference (never copied), plus oracle/ref_driver.cu in place of its GLUT main.cpp.
# Test infrastructure (second parity witness, same-box CUDA baseline).  Shims, all on the command line
# (SURVEY App. F): -DcutStatus=Status (undoes the author's local CUTLASS rename, readme.txt:3-5),
# -include thrust/execution_policy.h, an empty png.h, and this repository's HighFive-subset headers
# (the reference needs libhdf5, which exists on neither machine).
set -e
HERE="$(cd "$(dirname "$0")" && pwd)"
R=${REFERENCE_ROOT:-/root/reference}
OUT="$HERE/_ref"
[ -d "$R/src" ] || { echo "build_ref: $R/src not found, skipping"; exit 0; }
[ -f "$R/submodules/cutlass/include/cutlass/cutlass.h" ] || { echo "build_ref: vendored CUTLASS missing, skipping"; exit 0; }
mkdir -p "$OUT/obj"
NVCC=${NVCC:-nvcc}
FL="-std=c++14 -O2 -DNDEBUG -DcutStatus=Status -gencode arch=compute_100,code=sm_100 -w \
    -I$HERE/ref_shims \
    -I$R/src -I$R/src/layers -I$R/src/common -I$R/src/neuralUtils -I$R/submodules/cutlass/include \
    -I$HERE/../include"   # LAST: only <highfive/...> may resolve here; every other header is the reference's own
if [ "$OUT/ref_driver" -nt "$HERE/ref_driver.cu" ] && [ "$OUT/ref_driver" -nt "$HERE/build_ref.sh" ]; then
    echo "build_ref: $OUT/ref_driver is up to date"; exit 0
fi
$NVCC $FL -rdc=true -c "$R/src/layers/denseLayer.cu" -o "$OUT/obj/denseLayer.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/matrix.cu" -o "$OUT/obj/matrix.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/shape.cu" -o "$OUT/obj/shape.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/image.cu" -o "$OUT/obj/image.o" &
$NVCC $FL -rdc=true -include thrust/execution_policy.h -c "$R/src/volumeRender_kernel.cu" -o "$OUT/obj/volumeRender_kernel.o" &
$NVCC $FL -x cu -rdc=true -c "$R/src/neuralNetwork.cpp" -o "$OUT/obj/neuralNetwork.o" &
g++ -O2 -w -c "$R/src/common/lodepng.cpp" -o "$OUT/obj/lodepng.o" &
$NVCC $FL -rdc=true -c "$HERE/ref_driver.cu" -o "$OUT/obj/ref_driver.o" &
wait
$NVCC -gencode arch=compute_100,code=sm_100 -rdc=true "$OUT"/obj/*.o -o "$OUT/ref_driver"
echo "build_ref: built $OUT/ref_driver"
