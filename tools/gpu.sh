#!/bin/bash
# tools/gpu.sh [gpurun options] -- '<command>' : rebuild every in-tree artefact first (a stale .so travels
# silently otherwise), then hand over to gpurun.
set -e
cd "$(dirname "$0")/.."
make -C nbody_b200/csrc -j8 >/dev/null
make -C oracle liboracle.so >/dev/null
exec /usr/local/graft/bin/gpurun "$@"
