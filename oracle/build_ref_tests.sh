#!/bin/bash
# oracle/build_ref_tests.sh -- compile the reference's OWN test sources for the hot path, unmodified, where they lie
# under /root/reference/tests, against THIS repo's drop-in headers (include/nbody/*.hpp) and libnbody_b200.so, with
# the small GoogleTest / RapidCheck stand-ins of tests/cpp/shim (neither library is in this image), into
# oracle/_ref/ref_tests_on_b200 (git-ignored, travels to the GPU box; tests/test_gpu_reference_suite.py runs it).
# test_camera / test_color_mapping use glm types: tests/cpp/shim/glm maps them onto this repo's Vec3 / Mat4.
# Left out: test_app_cli / test_ui_panel / test_hdf5_io (out of scope).
# TEST INFRASTRUCTURE: the product never links this.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
OUT="$HERE/_ref"
[ -d "$REF/tests" ] || { echo "reference not present at $REF; keeping any prebuilt $OUT/ref_tests_on_b200"; exit 0; }
[ -f "$ROOT/nbody_b200/libnbody_b200.so" ] || { echo "libnbody_b200.so not built yet"; exit 1; }
mkdir -p "$OUT"
SRCS=""
for t in main force_calculation integrator barnes_hut spatial_hash particle_data serialization performance_observability validation \
         camera color_mapping; do
  SRCS="$SRCS $REF/tests/test_$t.cpp"
done
/usr/bin/g++ -std=c++20 -O1 -w -include "$ROOT/tests/cpp/shim/glm/glm.hpp" -I"$ROOT/tests/cpp/shim" -I"$ROOT/include" $SRCS -o "$OUT/ref_tests_on_b200" \
  -L"$ROOT/nbody_b200" -lnbody_b200 -Wl,-rpath,'$ORIGIN/../../nbody_b200' -lpthread
echo "built $OUT/ref_tests_on_b200"
