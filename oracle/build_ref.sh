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
if [ "$OUT/ref_driver" -nt "$HERE/ref_driver.cu" ] && [ "$OUT/ref_driver" -nt "$HERE/build_ref.sh" ] && \
   [ "$OUT/ref_sdf_probe" -nt "$HERE/ref_sdf_probe.cu" ] && [ "$OUT/ref_sdf_probe" -nt "$HERE/build_ref.sh" ] && \
   [ "$OUT/ref_driver_tanh" -nt "$HERE/build_ref.sh" ]; then
    echo "build_ref: $OUT/ref_driver, ref_sdf_probe, ref_driver_tanh are up to date"; exit 0
fi
$NVCC $FL -rdc=true -c "$R/src/layers/denseLayer.cu" -o "$OUT/obj/denseLayer.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/matrix.cu" -o "$OUT/obj/matrix.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/shape.cu" -o "$OUT/obj/shape.o" &
$NVCC $FL -rdc=true -c "$R/src/neuralUtils/image.cu" -o "$OUT/obj/image.o" &
$NVCC $FL -rdc=true -include thrust/execution_policy.h -c "$R/src/volumeRender_kernel.cu" -o "$OUT/obj/volumeRender_kernel.o" &
$NVCC $FL -x cu -rdc=true -c "$R/src/neuralNetwork.cpp" -o "$OUT/obj/neuralNetwork.o" &
g++ -O2 -w -c "$R/src/common/lodepng.cpp" -o "$OUT/obj/lodepng.o" &
$NVCC $FL -rdc=true -c "$HERE/ref_driver.cu" -o "$OUT/obj/ref_driver.o" &
# ref_sdf_probe: the reference's volumeRender_kernel.cu #included unmodified + one kernel that calls its analytic-SDF
# device functions (most are unreachable from sceneSDF at HEAD), oracle/ref_sdf_probe.cu
$NVCC $FL -rdc=true -include thrust/execution_policy.h -c "$HERE/ref_sdf_probe.cu" -o "$OUT/obj_probe.o" &
# ref_driver_tanh: the reference with sceneSDF's two early returns (src/volumeRender_kernel.cu:221-222) commented out, so
# that its own last line `return tanh(nSDF);` (:229, the variant the shipped .ppm renders were made with) is what runs.
# The two-line edit is applied by sed to a temporary copy that is deleted after compiling; nothing is stored.
TMPD="$(mktemp -d)"
sed -e 's|^\( *\)return manySphere(p, nSDF, true);|\1//return manySphere(p, nSDF, true);|' \
    -e 's|^\( *\)return manyCylinderCut(p, nSDF);|\1//return manyCylinderCut(p, nSDF);|' \
    "$R/src/volumeRender_kernel.cu" > "$TMPD/volumeRender_kernel_tanh.cu"
if [ "$(diff "$R/src/volumeRender_kernel.cu" "$TMPD/volumeRender_kernel_tanh.cu" | grep -c '^>')" != "2" ]; then
    echo "build_ref: the sceneSDF edit did not apply to exactly two lines"; rm -rf "$TMPD"; exit 1
fi
$NVCC $FL -rdc=true -include thrust/execution_policy.h -c "$TMPD/volumeRender_kernel_tanh.cu" -o "$OUT/obj_vrk_tanh.o" &
wait
rm -rf "$TMPD"
$NVCC -gencode arch=compute_100,code=sm_100 -rdc=true "$OUT"/obj/*.o -o "$OUT/ref_driver"
COMMON="$OUT/obj/denseLayer.o $OUT/obj/matrix.o $OUT/obj/shape.o $OUT/obj/image.o $OUT/obj/neuralNetwork.o $OUT/obj/lodepng.o"
$NVCC -gencode arch=compute_100,code=sm_100 -rdc=true $COMMON "$OUT/obj_probe.o" -o "$OUT/ref_sdf_probe"
$NVCC -gencode arch=compute_100,code=sm_100 -rdc=true $COMMON "$OUT/obj/ref_driver.o" "$OUT/obj_vrk_tanh.o" -o "$OUT/ref_driver_tanh"
echo "build_ref: built $OUT/ref_driver, $OUT/ref_sdf_probe, $OUT/ref_driver_tanh"
