mkdir -p gpurun_out; rm -f gpurun_out/paths9.jsonl
(timeout 900 python -m pytest tests/test_gpu_render.py tests/test_gpu_grid.py -m gpu -x -q > gpurun_out/pytest9.log 2>&1; echo "pytest exit $?" >> gpurun_out/pytest9.log); tail -12 gpurun_out/pytest9.log
python tools/bench_paths.py render 1048576 >> gpurun_out/paths9.jsonl 2>>gpurun_out/paths9.err
python tools/bench_paths.py render 16777216 >> gpurun_out/paths9.jsonl 2>>gpurun_out/paths9.err
python tools/bench_paths.py sh 10000000 >> gpurun_out/paths9.jsonl 2>>gpurun_out/paths9.err
cat gpurun_out/paths9.jsonl; tail -5 gpurun_out/paths9.err
# A/B on one box: the library as of commit 59d6739 against today's, direct N^2 1 Mi
for lib in build/ab/libnbody_b200_59d6739.so nbody_b200/libnbody_b200.so; do
NB_LIB_PATH=$PWD/$lib python - <<'PY'
import os, ctypes as C, numpy as np, torch
import nbody_b200 as nb
from nbody_b200 import _lib, ics
L=_lib.lib(); ctx=nb.get_context()
n=1<<20; st=ics.plummer(n,seed=42)
h=nb.ParticleDataManager.fromState(st); d=nb.ParticleData(); nb.ParticleDataManager.allocateDevice(d,n); nb.ParticleDataManager.copyToDevice(d,h)
for _ in range(2): _lib.check(L.nb_step_direct(ctx.handle,C.byref(d.c),1e-3,1.0,1e-4,None))
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(4): _lib.check(L.nb_step_direct(ctx.handle,C.byref(d.c),1e-3,1.0,1e-4,None))
e1.record(); torch.cuda.synchronize()
print(os.environ["NB_LIB_PATH"].split("/")[-1], "ms/step", e0.elapsed_time(e1)/4)
PY
done
