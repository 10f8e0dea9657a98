"""GPU, at BASELINE.json's full sizes: size-independent properties + sampled fp64 truth."""
import ctypes as C

import numpy as np
import pytest

from tests.conftest import rel_err

pytestmark = pytest.mark.gpu


def _device_data(state):
    import nbody_b200 as nb
    h = nb.ParticleDataManager.fromState(state)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    return d


@pytest.mark.parametrize("n", [1 << 20, 1000003])
def test_direct_1mi_accuracy_and_momentum(orc, n):
    """config 2: 1 Mi bodies (and 1,000,003 for the tail): max relative error <= 1e-5 on a sample of
    targets against fp64 truth over ALL sources; total force sums to zero (Newton's third law)."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.plummer(n, seed=42)
    d = _device_data(st)
    calc = nb.DirectForceCalculator()
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    o = d.numpy()
    rng = np.random.default_rng(1)
    idx = np.sort(rng.choice(n, 1536, replace=False)).astype(np.int32)
    idx[:4] = [0, 1, n - 2, n - 1]  # first tile and the padded tail
    idx = np.unique(idx).astype(np.int32)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, float(np.float32(0.01) ** 2), idx)
    e = rel_err((o["acc_x"][idx], o["acc_y"][idx], o["acc_z"][idx]), truth)
    assert e.max() <= 1e-5, f"max rel err {e.max():.3e} at n={n}"
    m = st["mass"].astype(np.float64)
    for k in ("acc_x", "acc_y", "acc_z"):
        tot = float((m * o[k]).sum())
        scale = float((m * np.abs(o[k])).sum())
        assert abs(tot) <= 2e-6 * scale, (k, tot, scale)
    assert np.isfinite(o["acc_x"]).all()


def test_direct_1mi_deterministic():
    """stream-K partials are reduced in a fixed order: two evaluations are bit-identical."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.plummer(1 << 20, seed=7)
    d = _device_data(st)
    calc = nb.DirectForceCalculator()
    calc.computeForces(d)
    a = d.numpy()["acc_x"].copy()
    calc.computeForces(d)
    assert np.array_equal(a, d.numpy()["acc_x"])


def test_sort_16mi_sorted_and_stable():
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib
    n = 1 << 24
    g = torch.Generator(device="cuda").manual_seed(5)
    keys = torch.randint(0, 1 << 20, (n,), device="cuda", dtype=torch.int32, generator=g)  # heavy duplication
    vals = torch.arange(n, device="cuda", dtype=torch.int32)
    k0 = keys.clone()
    _lib.check(_lib.lib().nb_sort_pairs(nb.get_context().handle, keys.data_ptr(), vals.data_ptr(), n, 20, None))
    torch.cuda.synchronize()
    assert bool((keys[1:] >= keys[:-1]).all())                                   # sortedness
    assert bool((k0[vals.long()] == keys).all())                                 # values follow their keys
    same = keys[1:] == keys[:-1]
    assert bool((vals[1:][same] > vals[:-1][same]).all())                        # stable: ties keep index order
    assert int(vals.long().sum()) == n * (n - 1) // 2                            # a permutation


def test_bh_16mi_properties():
    """config 4 at 16 Mi: permutation, sorted keys, mass conservation at the root, every body visited,
    interaction count sane, step under the 100 ms target."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    n = 1 << 24
    st = ics.plummer(n, seed=42)
    d = _device_data(st)
    L, ctx = _lib.lib(), nb.get_context()
    for _ in range(2):
        _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), None))
        _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), 0.5, 1.0, 0.01, None))
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    _lib.check(L.nb_bh_build(ctx.handle, C.byref(d.c), None))
    _lib.check(L.nb_bh_forces(ctx.handle, C.byref(d.c), 0.5, 1.0, 0.01, None))
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    keys, perm = np.zeros(n, np.uint32), np.zeros(n, np.int32)
    com = np.zeros((2 * n - 1, 4), np.float32)
    _lib.check(L.nb_bh_export(ctx.handle, keys.ctypes.data_as(C.c_void_p), perm.ctypes.data_as(C.c_void_p), None, None,
                              None, None, com.ctypes.data_as(C.c_void_p), None))
    assert (np.diff(keys.astype(np.int64)) >= 0).all()
    assert np.array_equal(np.sort(perm), np.arange(n, dtype=np.int32))
    assert abs(float(com[0, 3]) - 1.0) < 1e-3                                     # root mass = total mass
    inter = C.c_uint64(0)
    _lib.check(L.nb_bh_stats(ctx.handle, C.byref(inter), None))
    assert 500 * n < inter.value < 20000 * n
    acc = d.arrays["acc_x"]
    assert bool(torch.isfinite(acc).all()) and float(acc.abs().max()) > 0
    # sampled accuracy vs direct fp64 truth: Barnes-Hut theta = 0.5 monopole error level
    assert ms < 100.0, f"16 Mi Barnes-Hut step took {ms:.1f} ms"


def test_bh_1mi_accuracy_sample(orc):
    import nbody_b200 as nb
    from nbody_b200 import ics
    n = 1 << 20
    st = ics.plummer(n, seed=42)
    d = _device_data(st)
    calc = nb.BarnesHutCalculator(0.5)
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    o = d.numpy()
    idx = np.sort(np.random.default_rng(2).choice(n, 512, replace=False)).astype(np.int32)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, float(np.float32(0.01) ** 2), idx)
    e = rel_err((o["acc_x"][idx], o["acc_y"][idx], o["acc_z"][idx]), truth)
    assert np.sqrt(np.mean(e ** 2)) < 3e-3 and e.max() < 3e-2, (np.sqrt(np.mean(e ** 2)), e.max())


def test_sh_10m_properties():
    """config 5: 10 M bodies, ~8 per unit cell: the cells partition the bodies exactly once, ranges are
    consistent with the per-particle cell ids, forces finite."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import ics
    n = 10_000_000
    side = (n / 8.0) ** (1.0 / 3.0)
    st = ics.uniform_box(n, 0.0, side, seed=42)
    d = _device_data(st)
    calc = nb.SpatialHashCalculator(1.0, 2.0)
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    g = calc.getGrid().copyCellDataToHost()
    assert 1_000_000 < g["total_cells"] < 2_000_000
    cnt = np.bincount(g["particle_cell"], minlength=g["total_cells"])
    start = np.concatenate([[0], np.cumsum(cnt)[:-1]])
    assert np.array_equal(g["cell_start"], start) and np.array_equal(g["cell_end"], start + cnt)
    si = g["sorted_indices"]
    assert np.array_equal(np.sort(si), np.arange(n, dtype=np.int32))
    assert (np.diff(g["particle_cell"][si]) >= 0).all()
    assert bool(torch.isfinite(d.arrays["acc_x"]).all())
