#!/bin/bash
# integration/build_dropin.sh -- proof of the drop-in claim: the reference's OWN benchmark program
# (benchmarks/benchmark_main.cpp) and facade (src/core/particle_system.cpp), compiled from the
# unmodified sources under /root/reference, linked against integration/b200_backend.cpp +
# libnbody_b200.so instead of the reference's four hot-path .cu files.  Output:
# oracle/_ref/ref_bench_on_b200 (git-ignored, travels to the GPU box).  Test infrastructure only.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"; ROOT="$(dirname "$HERE")"
OUT="$ROOT/oracle/_ref"
[ -d "$REF/src/core" ] || { echo "reference not present at $REF; keeping any prebuilt binary"; exit 0; }
mkdir -p "$OUT/obj2"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
F="-std=c++17 -O3 -arch=sm_100 -DNDEBUG -DNBODY_WITH_CUDA=1 -DNBODY_WITH_RENDERING=0 -DNBODY_WITH_UI=0 -DNBODY_WITH_HDF5=0 -DNBODY_ENABLE_PROFILING=1 -I$REF/include -I$ROOT/include -Xcompiler -fPIC -ccbin /usr/bin/g++ -w"
# host .cpp go through `nvcc -x cu` only to dodge the reference's int3 redefinition
for f in src/core/particle_system src/core/app_cli src/utils/error_handling src/utils/performance_observability \
         src/utils/serialization src/utils/simulation_state benchmarks/benchmark_main; do
  $NVCC $F -x cu -c "$REF/$f.cpp" -o "$OUT/obj2/$(basename $f).o" &
done
$NVCC $F -c "$REF/src/cuda/particle_init.cu" -o "$OUT/obj2/particle_init.o" &
$NVCC $F -x cu -c "$HERE/b200_backend.cpp" -o "$OUT/obj2/b200_backend.o" &
wait
$NVCC -arch=sm_100 -o "$OUT/ref_bench_on_b200" "$OUT"/obj2/*.o -L"$ROOT/nbody_b200" -lnbody_b200 \
  -Xlinker -rpath -Xlinker '$ORIGIN/../../nbody_b200' -ccbin /usr/bin/g++
rm -rf "$OUT/obj2"
echo "built $OUT/ref_bench_on_b200"
