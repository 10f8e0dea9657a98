#!/usr/bin/env python
"""Per-stage timings of the Barnes-Hut and spatial-hash paths (CUDA events), configs 4 and 5 of
BASELINE.json.  usage: tools/bench_paths.py bh 1048576 [plummer|sphere]  |  sh 10000000"""
import ctypes as C
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import nbody_b200 as nb  # noqa: E402
from nbody_b200 import _lib, ics  # noqa: E402


def timed(fn, reps=5, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    kind, n = sys.argv[1], int(sys.argv[2])
    L, ctx = _lib.lib(), nb.get_context()
    if kind == "bh":
        ic = sys.argv[3] if len(sys.argv) > 3 else "plummer"
        st = ics.plummer(n, seed=42) if ic == "plummer" else ics.uniform_sphere(n, 1.0, 1.0 / n, seed=42)
        h = nb.ParticleDataManager.fromState(st)
        d = nb.ParticleData()
        nb.ParticleDataManager.allocateDevice(d, n)
        nb.ParticleDataManager.copyToDevice(d, h)
        s = None
        build = timed(lambda: _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), s)))
        force = timed(lambda: _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), 0.5, 1.0, 0.01, s)))
        _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), s))
        _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), 0.5, 1.0, 0.01, s))
        inter, cnt = C.c_uint64(0), C.c_int(0)
        _lib.check(L.nb_bh_stats(ctx.handle, C.byref(inter), C.byref(cnt)))
        deep, cap = C.c_int(0), C.c_int(0)
        _lib.check(L.nb_bh_max_stack(ctx.handle, C.byref(deep), C.byref(cap)))
        print(json.dumps(dict(path="barnes_hut", n=n, ic=ic, theta=0.5, group=os.environ.get("NB_BH_GROUP", "64"),
                              stack_deepest=deep.value, stack_cap=cap.value,
                              super=os.environ.get("NB_BH_SUPER", "default"), build_ms=build, force_ms=force,
                              step_ms=build + force, interactions=inter.value, per_body=inter.value / n,
                              g_interactions_per_s=inter.value / force / 1e6)))
    elif kind == "energy":
        # potential-energy diagnostic (O(N^2)) and the one-pass moments, wall clock around the synchronising calls
        import time
        st = ics.plummer(n, seed=42)
        h = nb.ParticleDataManager.fromState(st)
        d = nb.ParticleData()
        nb.ParticleDataManager.allocateDevice(d, n)
        nb.ParticleDataManager.copyToDevice(d, h)
        integ = nb.Integrator()
        integ.computePotentialEnergy(d, 1.0, 0.01)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            pe = integ.computePotentialEnergy(d, 1.0, 0.01)
        pe_ms = (time.perf_counter() - t0) / reps * 1e3
        t0 = time.perf_counter()
        for _ in range(reps):
            dg = integ.computeDiagnostics(d, 1.0, 0.01, with_potential=False)
        mo_ms = (time.perf_counter() - t0) / reps * 1e3
        print(json.dumps(dict(path="energy", n=n, potential=pe, potential_ms=pe_ms,
                              g_pairs_per_s=0.5 * n * (n - 1) / pe_ms / 1e6, moments_ms=mo_ms, kinetic=dg["kinetic"])))
    elif kind == "graph":
        # steps/s of ParticleSystem.update, eager launches against CUDA-graph replay (launch-bound sizes)
        method = nb.ForceMethod.BARNES_HUT if (len(sys.argv) > 3 and sys.argv[3] == "bh") else nb.ForceMethod.DIRECT_N2
        st = ics.plummer(n, seed=42)
        out = {}
        for mode in ("eager", "graph"):
            s = nb.ParticleSystem()
            s.config_ = nb.SimulationConfig(particle_count=n, softening=0.01, dt=1e-3, force_method=method)
            s.setState(nb.SimulationState(n, 0.0, 1e-3, 1.0, 0.01, method,
                                          {k: st[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")}))
            s.setUseCudaGraph(mode == "graph")
            for _ in range(3):
                s.update(1e-3)
            out[mode] = timed(lambda: s.update(1e-3), reps=100, warm=5)
            out[mode + "_nodes"] = s.graphNodeCount()
        print(json.dumps(dict(path="step_graph", method=method.name, n=n, eager_ms=out["eager"], graph_ms=out["graph"],
                              graph_nodes=out["graph_nodes"], eager_steps_per_s=1e3 / out["eager"],
                              graph_steps_per_s=1e3 / out["graph"], speedup=out["eager"] / out["graph"])))
    elif kind == "render":
        # default camera and uniforms of the reference's Renderer; Plummer sphere scaled into view
        w, h = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (1280, 720)
        st = ics.plummer(n, seed=42)
        for k in ("pos_x", "pos_y", "pos_z"):
            st[k] = (st[k] * np.float32(8.0)).astype(np.float32)
        hd = nb.ParticleDataManager.fromState(st)
        d = nb.ParticleData()
        nb.ParticleDataManager.allocateDevice(d, n)
        nb.ParticleDataManager.copyToDevice(d, hd)
        r = nb.Renderer()
        r.initialize(w, h)
        ms = timed(lambda: r.render(d))
        frags = r.fragments()
        img = r.readPixels()
        r.savePPM(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out",
                               f"render_{n}.ppm")) if os.path.isdir("gpurun_out") and n <= (1 << 20) else None
        print(json.dumps(dict(path="splat_render", n=n, width=w, height=h, ms=ms, fragments=frags,
                              lit_pixels=int((img != img[0, 0]).any(axis=2).sum()), gbodies_per_s=n / ms / 1e6)))
    elif kind == "sort":
        # 30-bit keys in random order (what the Morton sort sees), restored before every repetition
        bits = int(sys.argv[3]) if len(sys.argv) > 3 else 30
        g = torch.Generator(device="cuda").manual_seed(42)
        k0 = torch.randint(0, 1 << bits, (n,), device="cuda", dtype=torch.int64, generator=g).to(torch.int32)
        v0 = torch.arange(n, device="cuda", dtype=torch.int32)
        k, v = k0.clone(), v0.clone()
        ms = []
        for rep in range(7):
            k.copy_(k0); v.copy_(v0)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            _lib.check(L.nb_sort_pairs(ctx.handle, k.data_ptr(), v.data_ptr(), n, bits, None))
            e1.record()
            torch.cuda.synchronize()
            ms.append(e0.elapsed_time(e1))
        best = min(ms[2:])
        passes = (bits + 7) // 8
        algo = (4 + 16 * passes) * n
        assert bool((k[1:] >= k[:-1]).all())
        print(json.dumps(dict(path="radix_sort_pairs", n=n, bits=bits, passes=passes, ms=best, ms_all=ms,
                              warps=os.environ.get("NB_SORT_WARPS", "16"), algorithmic_bytes=algo,
                              gbs=algo / best / 1e6, gkeys_per_s=n / best / 1e6)))
    else:
        side = (n / 8.0) ** (1.0 / 3.0)     # ~8 bodies per unit cell
        st = ics.uniform_box(n, 0.0, side, seed=42)
        h = nb.ParticleDataManager.fromState(st)
        d = nb.ParticleData()
        nb.ParticleDataManager.allocateDevice(d, n)
        nb.ParticleDataManager.copyToDevice(d, h)
        build = timed(lambda: _lib.check(L.nb_sh_build(ctx.handle, C.byref(d.c), 1.0, None)))
        force = timed(lambda: _lib.check(L.nb_sh_forces(ctx.handle, C.byref(d.c), 2.0, 1.0, 0.01, None)))
        t, pairs = C.c_int(0), C.c_uint64(0)
        L.nb_sh_total_cells(ctx.handle, C.byref(t))
        _lib.check(L.nb_sh_stats(ctx.handle, C.byref(pairs)))
        print(json.dumps(dict(path="spatial_hash", n=n, side=side, cells=t.value, build_ms=build, force_ms=force,
                              step_ms=build + force, build_gbs=136.0 * n / build / 1e6, pairs_in_cutoff=pairs.value,
                              pairs_per_body=pairs.value / n, g_pairs_per_s=pairs.value / force / 1e6)))


if __name__ == "__main__":
    main()
