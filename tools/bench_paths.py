#!/usr/bin/env python
"""Per-stage timings of the Barnes-Hut, spatial-hash and sort paths (CUDA events), configs 4 and 5 of
BASELINE.json.  bench.py imports the functions below for its `paths` block; stand-alone:
  tools/bench_paths.py bh 1048576 [plummer|sphere] | sh 10000000 | sort 16777216 [bits] | energy N | graph N [bh] | render N

Every HBM-side stage is reported with its ALGORITHMIC bytes (what the stage has to move with our layouts,
DESIGN.md 7.1), the achieved GB/s over its CUDA-event time, and the fraction of the measured copy bandwidth
(MEASURED_PEAKS.json hbm_gbs, else the B200_PROFILING.md fallback)."""
import ctypes as C
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import nbody_b200 as nb  # noqa: E402
from nbody_b200 import _lib, ics  # noqa: E402

HBM_FALLBACK_GBS = 6400.0  # B200_PROFILING.md: measured copy bandwidth to use when MEASURED_PEAKS.json is absent

# algorithmic bytes per body of each stage, with the layouts of DESIGN.md sec. 2
ALGO_BYTES = {
    "bh.bbox": 12,            # x, y, z read once
    "bh.morton": 16 + 4 + 16,  # x,y,z,m read; key and packed {x,y,z,m} written (the sort generates the identity values)
    # keys + perm read, packed float4 gathered, leaf record written; per internal node {left,right,prefix} + centre of mass
    "bh.tree": 4 + 4 + 16 + 16 + 16 + 16,
    # node record + both children's centres of mass + their levels read, one 48-byte pair record written per node
    "bh.pairs": 16 + 32 + 8 + 48,
    "sh.bbox": 12,
    "sh.assign": 16 + 4 + 4 + 4 + 16,  # x,y,z,m read; particle_cell, key, identity value and packed written
    "sh.ranges_gather": 4 + 4 + 16 + 16,  # sorted key + index read, packed gathered, cell-ordered float4 written (+ cell ranges)
    "sh.pair_list": 2 * 8 + 4,     # sorted key + cell_start read twice (flag pass, fill pass), pair slot written
}


def hbm_peak_gbs():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return HBM_FALLBACK_GBS, "B200_PROFILING.md fallback"


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


def sort_passes(bits):
    return int(_lib.lib().nb_sort_passes(bits))


def sort_algo_bytes(n, bits, identity_values=False):
    """keys read once for all digit histograms, then 16 B per pair and pass (8 read + 8 written); a sort whose
    values are the identity (the tree and grid builds) does not read them in the first pass."""
    return (4 + 16 * sort_passes(bits) - (4 if identity_values else 0)) * n


def stages(fn, reps, n, sort_bits=None, identity_values=False):
    """Run fn reps times under the library's stage timers; returns {stage: {ms, algorithmic_bytes, gbs, frac}}."""
    L, ctx = _lib.lib(), nb.get_context()
    peak, _ = hbm_peak_gbs()
    buf = C.create_string_buffer(1 << 16)
    torch.cuda.synchronize()
    _lib.check(L.nb_profile_stages_enable(ctx.handle, 1))
    _lib.check(L.nb_profile_stages_read(ctx.handle, buf, len(buf)))  # drop anything older
    for _ in range(reps):
        fn()
    _lib.check(L.nb_profile_stages_read(ctx.handle, buf, len(buf)))
    _lib.check(L.nb_profile_stages_enable(ctx.handle, 0))
    out = {}
    for ln in buf.value.decode().splitlines():
        name, ms, cnt = ln.split()
        ms = float(ms) / reps
        rec = {"ms": ms, "launches_per_call": int(cnt) / reps}
        b = None
        if name.endswith(".sort") and sort_bits:
            b = sort_algo_bytes(n, sort_bits, identity_values)
        elif name in ALGO_BYTES:
            b = ALGO_BYTES[name] * n
        if b is not None and ms > 0:
            rec.update(algorithmic_bytes=int(b), achieved_gbs=b / ms / 1e6, frac=b / ms / 1e6 / peak)
        out[name] = rec
    return out


def device_data(st):
    h = nb.ParticleDataManager.fromState(st)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    return d


def bench_bh(n, ic="plummer", theta=0.5, reps=5):
    L, ctx = _lib.lib(), nb.get_context()
    st = ics.plummer(n, seed=42) if ic == "plummer" else ics.uniform_sphere(n, 1.0, 1.0 / n, seed=42)
    d = device_data(st)
    s = None
    build = timed(lambda: _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), s)), reps)
    force = timed(lambda: _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), theta, 1.0, 0.01, s)), reps, warm=1)

    def step():
        _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), s))
        _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), theta, 1.0, 0.01, s))
    step_ms = timed(step, reps, warm=1)
    stg = stages(step, 3, n, sort_bits=30, identity_values=True)
    inter, cnt = C.c_uint64(0), C.c_int(0)
    _lib.check(L.nb_bh_stats(ctx.handle, C.byref(inter), C.byref(cnt)))
    deep, cap = C.c_int(0), C.c_int(0)
    _lib.check(L.nb_bh_max_stack(ctx.handle, C.byref(deep), C.byref(cap)))
    peak, src = hbm_peak_gbs()
    build_bytes = sum(r.get("algorithmic_bytes", 0) for k, r in stg.items() if k != "bh.traversal")
    return dict(path="barnes_hut", n=n, ic=ic, theta=theta, build_ms=build, force_ms=force, step_ms=step_ms,
                build_algorithmic_bytes=int(build_bytes), build_gbs=build_bytes / build / 1e6,
                build_frac=build_bytes / build / 1e6 / peak, hbm_peak_gbs=peak, hbm_peak_source=src,
                interactions=inter.value, interactions_per_body=inter.value / n,
                g_interactions_per_s=inter.value / force / 1e6,
                tflops_20flop=inter.value * 20.0 / force / 1e9, stack_deepest=deep.value, stack_cap=cap.value,
                stages=stg, timing="CUDA events, mean of %d calls after warm-up; stages under the library's stage "
                                   "timers (events between the kernels of one call)" % reps)


def bench_sh(n, reps=5):
    L, ctx = _lib.lib(), nb.get_context()
    side = (n / 8.0) ** (1.0 / 3.0)     # ~8 bodies per unit cell
    st = ics.uniform_box(n, 0.0, side, seed=42)
    d = device_data(st)
    build = timed(lambda: _lib.check(L.nb_sh_build(ctx.handle, C.byref(d.c), 1.0, None)), reps)
    force = timed(lambda: _lib.check(L.nb_sh_forces(ctx.handle, C.byref(d.c), 2.0, 1.0, 0.01, None)), reps)
    t, pairs = C.c_int(0), C.c_uint64(0)
    L.nb_sh_total_cells(ctx.handle, C.byref(t))
    _lib.check(L.nb_sh_stats(ctx.handle, C.byref(pairs)))
    bits = max(1, int(np.ceil(np.log2(max(t.value, 2)))))

    def step():
        _lib.check(L.nb_sh_build(ctx.handle, C.byref(d.c), 1.0, None))
        _lib.check(L.nb_sh_forces(ctx.handle, C.byref(d.c), 2.0, 1.0, 0.01, None))
    stg = stages(step, 3, n, sort_bits=bits, identity_values=False)
    peak, src = hbm_peak_gbs()
    build_bytes = sum(r.get("algorithmic_bytes", 0) for k, r in stg.items() if k != "sh.force")
    return dict(path="spatial_hash", n=n, side=side, cell_size=1.0, cutoff=2.0, cells=t.value, key_bits=bits,
                build_ms=build, force_ms=force, step_ms=build + force,
                build_algorithmic_bytes=int(build_bytes), build_gbs=build_bytes / build / 1e6,
                build_frac=build_bytes / build / 1e6 / peak, hbm_peak_gbs=peak, hbm_peak_source=src,
                pairs_in_cutoff=pairs.value, pairs_per_body=pairs.value / n, g_pairs_per_s=pairs.value / force / 1e6,
                stages=stg)


def bench_sort(n, bits=30, reps=7):
    """(key, value) pairs with `bits`-bit keys in random order (what the Morton sort sees), restored before every
    repetition; nb_sort_pairs sorts in place."""
    L, ctx = _lib.lib(), nb.get_context()
    g = torch.Generator(device="cuda").manual_seed(42)
    k0 = torch.randint(0, 1 << bits, (n,), device="cuda", dtype=torch.int64, generator=g).to(torch.int32)
    v0 = torch.arange(n, device="cuda", dtype=torch.int32)
    k, v = k0.clone(), v0.clone()
    ms = []
    for rep in range(reps):
        k.copy_(k0); v.copy_(v0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.check(L.nb_sort_pairs(ctx.handle, k.data_ptr(), v.data_ptr(), n, bits, None))
        e1.record()
        torch.cuda.synchronize()
        ms.append(e0.elapsed_time(e1))
    best, mean = min(ms[2:]), float(np.mean(ms[2:]))
    passes = sort_passes(bits)
    algo = sort_algo_bytes(n, bits) + (16 * n if passes % 2 else 0)  # in place: odd pass counts copy back
    assert bool((k[1:] >= k[:-1]).all())
    assert bool((k0[v.long()] == k).all())
    peak, src = hbm_peak_gbs()
    return dict(path="radix_sort_pairs", n=n, bits=bits, passes=passes, ms=mean, ms_best=best, ms_all=ms,
                algorithmic_bytes=int(algo), achieved_gbs=algo / mean / 1e6, frac=algo / mean / 1e6 / peak,
                hbm_peak_gbs=peak, hbm_peak_source=src, gkeys_per_s=n / mean / 1e6,
                what="nb_sort_pairs in place: keys read once for the histograms, 16 B per pair and pass, and after an "
                     "odd number of passes 16 B per pair for the copy back into the caller's arrays (the tree and "
                     "grid builds take the result from the scratch buffer instead: see paths.*.stages['*.sort'])")


def main():
    kind, n = sys.argv[1], int(sys.argv[2])
    if kind == "bh":
        print(json.dumps(bench_bh(n, sys.argv[3] if len(sys.argv) > 3 else "plummer")))
    elif kind == "sh":
        print(json.dumps(bench_sh(n)))
    elif kind == "sort":
        print(json.dumps(bench_sort(n, int(sys.argv[3]) if len(sys.argv) > 3 else 30)))
    elif kind == "energy":
        # potential-energy diagnostic (O(N^2)) and the one-pass moments, wall clock around the synchronising calls
        import time
        st = ics.plummer(n, seed=42)
        d = device_data(st)
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
        d = device_data(st)
        r = nb.Renderer()
        r.initialize(w, h)
        ms = timed(lambda: r.render(d))
        frags = r.fragments()
        img = r.readPixels()
        if os.path.isdir(os.path.join(ROOT, "gpurun_out")) and n <= (1 << 20):
            r.savePPM(os.path.join(ROOT, "gpurun_out", f"render_{n}.ppm"))
        print(json.dumps(dict(path="splat_render", n=n, width=w, height=h, ms=ms, fragments=frags,
                              lit_pixels=int((img != img[0, 0]).any(axis=2).sum()), gbodies_per_s=n / ms / 1e6)))
    else:
        raise SystemExit(__doc__)


if __name__ == "__main__":
    main()
