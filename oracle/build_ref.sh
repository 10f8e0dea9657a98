#!/bin/bash
# oracle/build_ref.sh -- compile the UNMODIFIED reference CUDA sources where they lie under
# /root/reference for sm_100 into oracle/_ref/libnbody_ref.so (git-ignored, travels to the GPU box).
# The reference's own CMake CUDA configuration does not compile (int3 redefinition in host TUs,
# incomplete ForceCalculator in integrator.cu); the flag-only workarounds are:
#   * host .cpp compiled with `nvcc -x cu`      (cuda_runtime.h's int3 is used instead of the shim's)
#   * `-include nbody/force_calculator.hpp` for integrator.cu
#   * -rdc=true (force_spatial_hash.cu references the bbox kernel defined in force_barnes_hut.cu)
# TEST INFRASTRUCTURE: the product never links this.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
[ -d "$REF/src/cuda" ] || { echo "reference not present at $REF; keeping any prebuilt $OUT"; exit 0; }
mkdir -p "$OUT/obj"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
F="-std=c++17 -O3 -arch=sm_100 -rdc=true -DNDEBUG -DNBODY_WITH_CUDA=1 -DNBODY_WITH_RENDERING=0 -DNBODY_WITH_UI=0 -DNBODY_WITH_HDF5=0 -I$REF/include -Xcompiler -fPIC -ccbin /usr/bin/g++ -w"
for f in force_direct force_barnes_hut force_spatial_hash particle_init; do
  $NVCC $F -c "$REF/src/cuda/$f.cu" -o "$OUT/obj/$f.o" &
done
$NVCC $F -include nbody/force_calculator.hpp -c "$REF/src/cuda/integrator.cu" -o "$OUT/obj/integrator.o" &
$NVCC $F -x cu -c "$REF/src/utils/error_handling.cpp" -o "$OUT/obj/error_handling.o" &
$NVCC $F -c "$HERE/ref_shim.cu" -o "$OUT/obj/ref_shim.o" &
wait
$NVCC -arch=sm_100 -shared -o "$OUT/libnbody_ref.so" "$OUT"/obj/*.o -ccbin /usr/bin/g++
rm -rf "$OUT/obj"
echo "built $OUT/libnbody_ref.so"
