#!/bin/bash
# oracle/build_ref_host.sh -- compile the reference's UNMODIFIED host-side checkpoint and benchmark-record
# sources where they lie under /root/reference (plain g++, no CUDA, no CMake) together with
# oracle/ref_host_tool.cpp into oracle/_ref/ref_host_tool (git-ignored, travels to the GPU box).
# TEST INFRASTRUCTURE: the product never links this.
set -euo pipefail
REF=${REF:-/root/reference}
HERE="$(cd "$(dirname "$0")" && pwd)"
OUT="$HERE/_ref"
[ -d "$REF/src/utils" ] || { echo "reference not present at $REF; keeping any prebuilt $OUT/ref_host_tool"; exit 0; }
mkdir -p "$OUT"
/usr/bin/g++ -std=c++17 -O1 -w -I"$REF/include" -DNBODY_WITH_CUDA=0 -DNBODY_WITH_RENDERING=0 -DNBODY_WITH_UI=0 \
  -DNBODY_WITH_HDF5=0 "$HERE/ref_host_tool.cpp" "$REF/src/utils/serialization.cpp" "$REF/src/utils/simulation_state.cpp" \
  "$REF/src/utils/error_handling.cpp" "$REF/src/utils/performance_observability.cpp" -o "$OUT/ref_host_tool"
echo "built $OUT/ref_host_tool"
