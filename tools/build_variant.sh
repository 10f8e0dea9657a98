#!/bin/bash
# tools/build_variant.sh <tag> <file.cu> [nvcc -D flags...] : a second libnbody_b200_<tag>.so whose <file.cu> (direct.cu, sh.cu,
# bh.cu, ...) is compiled with the given build knobs (NB_DIRECT_UNROLL, NB_DIRECT_THREADS_PER_SM, NB_SH_MIN_BLOCKS, ...),
# every other object shared with the product build.  VARIANT_SRC=<path> compiles another revision of that file instead
# (e.g. `git show HEAD~1:nbody_b200/csrc/direct.cu > /tmp/d.cu`).  Time the result with tools/ab_direct.sh or
# NB_LIB_PATH=$PWD/nbody_b200/libnbody_b200_<tag>.so python tools/bench_paths.py ...; the tests honour NB_LIB_PATH too.
set -e
cd "$(dirname "$0")/../nbody_b200/csrc"
tag=$1; file=$2; shift 2
base=${file%.cu}
make -j8 >/dev/null
mkdir -p ../../build/var
src=${VARIANT_SRC:-$file}
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -ccbin /usr/bin/g++ \
  -I. "$@" -c "$src" -o ../../build/var/${base}_$tag.o
objs=$(ls ../../build/obj/*.o | grep -v "/$base.o")
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libnbody_b200_$tag.so $objs ../../build/var/${base}_$tag.o -ccbin /usr/bin/g++
echo "built nbody_b200/libnbody_b200_$tag.so"
