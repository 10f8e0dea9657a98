#!/bin/bash
# A/B of libnbody_b200 builds on ONE box: ms/step of nb_step_direct at 1 Mi bodies for each .so given
# AB_MIXED=1: one mass in every 512 differs in its last bit (every source tile takes the general pair loop)
for lib in "$@"; do
NB_LIB_PATH=$PWD/$lib python - <<'PY'
import os, ctypes as C, torch
import nbody_b200 as nb
from nbody_b200 import _lib, ics
L=_lib.lib(); ctx=nb.get_context()
n=1<<20; st=ics.plummer(n,seed=42)
import numpy as np
if os.environ.get('AB_MIXED'): st['mass'].view(np.uint32)[::512]^=1
h=nb.ParticleDataManager.fromState(st); d=nb.ParticleData(); nb.ParticleDataManager.allocateDevice(d,n); nb.ParticleDataManager.copyToDevice(d,h)
for _ in range(2): _lib.check(L.nb_step_direct(ctx.handle,C.byref(d.c),1e-3,1.0,1e-4,None))
torch.cuda.synchronize()
e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(4): _lib.check(L.nb_step_direct(ctx.handle,C.byref(d.c),1e-3,1.0,1e-4,None))
e1.record(); torch.cuda.synchronize()
print(os.environ["NB_LIB_PATH"].split("/")[-1], "mixed" if os.environ.get("AB_MIXED") else "equal", "ms/step %.3f" % (e0.elapsed_time(e1)/4), flush=True)
PY
done
