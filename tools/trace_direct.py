#!/usr/bin/env python
"""Experiment (needs a -DNB_DIRECT_TRACE build given through NB_LIB_PATH): which SM each CTA of the direct
force kernel ran on and when, for the 1 Mi step; prints CTAs per SM and the start/end spread."""
import ctypes as C
import os
import sys

import numpy as np
import torch

import nbody_b200 as nb
from nbody_b200 import _lib, ics

L = _lib.lib()
ctx = nb.get_context()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
st = ics.plummer(n, seed=42)
h = nb.ParticleDataManager.fromState(st)
d = nb.ParticleData()
nb.ParticleDataManager.allocateDevice(d, n)
nb.ParticleDataManager.copyToDevice(d, h)
for _ in range(2):
    _lib.check(L.nb_step_direct(ctx.handle, C.byref(d.c), 1e-3, 1.0, 1e-4, None))
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
_lib.check(L.nb_step_direct(ctx.handle, C.byref(d.c), 1e-3, 1.0, 1e-4, None))
e1.record()
torch.cuda.synchronize()
g = int(os.environ.get("NB_DIRECT_OCC", "5")) * 148
raw = C.CDLL(os.environ["NB_LIB_PATH"])
buf = np.zeros(3 * g, np.uint64)
rc = raw.nb_debug_direct_trace(buf.ctypes.data_as(C.c_void_p), g)
assert rc == 0, rc
t = buf.reshape(g, 3)
sm, t0, t1 = t[:, 0].astype(int), t[:, 1].astype(np.int64), t[:, 2].astype(np.int64)
base = t0.min()
per_sm = np.bincount(sm, minlength=148)
print(f"occ={g // 148} n={n} step {e0.elapsed_time(e1):.3f} ms; CTAs per SM: min {per_sm.min()} max {per_sm.max()} "
      f"hist {np.bincount(per_sm).tolist()}")
print(f"start spread: {(t0 - base).max() / 1e6:.3f} ms; CTAs starting after 1 ms: {int(((t0 - base) > 1e6).sum())}")
dur = (t1 - t0) / 1e6
print(f"duration ms: min {dur.min():.2f} median {np.median(dur):.2f} max {dur.max():.2f}; last end {(t1 - base).max() / 1e6:.2f} ms; "
      f"first end {(t1 - base).min() / 1e6:.2f} ms")
late = np.argsort(t0)[-8:]
print("latest starters (cta, sm, start ms, dur ms):", [(int(c), int(sm[c]), round(float(t0[c] - base) / 1e6, 2), round(float(dur[c]), 2)) for c in late])
