#!/bin/bash
# tools/build_direct_variant.sh <tag> [nvcc -D flags...] : a second libnbody_b200_<tag>.so whose direct.cu is compiled with
# the given knobs (NB_DIRECT_UNROLL, NB_DIRECT_THREADS_PER_SM, ...), every other object shared with the product build.
# DIRECT_SRC=<file> compiles another revision of direct.cu (e.g. `git show HEAD~1:nbody_b200/csrc/direct.cu`).
set -e
cd "$(dirname "$0")/../nbody_b200/csrc"
tag=$1; shift
make -j8 >/dev/null
mkdir -p ../../build/var
src=${DIRECT_SRC:-direct.cu}
/usr/local/cuda/bin/nvcc -O3 -std=c++17 -lineinfo -gencode arch=compute_100a,code=sm_100a -Xcompiler -fPIC -ccbin /usr/bin/g++ \
  -I. "$@" -c "$src" -o ../../build/var/direct_$tag.o
objs=$(ls ../../build/obj/*.o | grep -v /direct.o)
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libnbody_b200_$tag.so $objs ../../build/var/direct_$tag.o -ccbin /usr/bin/g++
cuobjdump -res-usage ../libnbody_b200_$tag.so 2>/dev/null | grep -A1 "direct_force_kernelILi256ELi2ELi1024ELb0" | grep -o "REG:[0-9]* STACK:[0-9]*" | sed "s/^/$tag: /"
