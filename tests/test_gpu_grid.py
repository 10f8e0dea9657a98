"""GPU parity: spatial-hash grid (bit-exact cell indices / ranges) and 27-cell forces vs the oracle."""
import numpy as np
import pytest

from tests.conftest import rel_err

pytestmark = pytest.mark.gpu


def _run(state, cell=1.0, cutoff=2.0, eps=0.01, G=1.0):
    import nbody_b200 as nb
    h = nb.ParticleDataManager.fromState(state)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    calc = nb.SpatialHashCalculator(cell, cutoff)
    calc.setSofteningParameter(eps)
    calc.setGravitationalConstant(G)
    calc.computeForces(d)
    return d, calc


@pytest.mark.parametrize("n,cell,kind", [(1, 1.0, "box"), (2, 1.0, "box"), (1000, 1.0, "box"), (5000, 2.0, "sphere"),
                                         (50000, 0.5, "box"), (200000, 1.0, "box"), (30000, 0.37, "plummer"),
                                         # few bodies on a fine grid: 9.3 M cells, almost all empty -- long runs of empty
                                         # cells between occupied ones (the gap list of sh_ranges_gather / sh_fill_gaps)
                                         (300, 0.095, "box"), (3, 0.1, "box")])
def test_grid_bit_exact(orc, n, cell, kind):
    from nbody_b200 import ics
    if kind == "sphere":
        st = ics.uniform_sphere(n, 10.0, seed=21)
    elif kind == "plummer":
        st = ics.plummer(n, seed=21)
    else:
        st = ics.uniform_box(n, -10, 10, seed=21)
    d, calc = _run(st, cell, 2 * cell)
    g = calc.getGrid().copyCellDataToHost()
    ref = orc.grid_build(st["pos_x"], st["pos_y"], st["pos_z"], cell)
    assert np.array_equal(g["dims"], ref["dims"])
    assert np.array_equal(g["bbox"], ref["bbox"])
    assert g["total_cells"] == ref["total_cells"]
    assert np.array_equal(g["particle_cell"], ref["particle_cell"])
    assert np.array_equal(g["cell_start"], ref["cell_start"])
    assert np.array_equal(g["cell_end"], ref["cell_end"])
    # our order inside a cell is ascending body index = the oracle's representative
    assert np.array_equal(g["sorted_indices"], ref["sorted_indices"])
    # cells partition all bodies exactly once (tests/test_spatial_hash.cpp:89-130)
    assert np.array_equal(np.sort(g["sorted_indices"]), np.arange(n))
    assert (g["cell_end"] - g["cell_start"]).sum() == n


@pytest.mark.parametrize("n,cell,cutoff", [(2, 1.0, 2.0), (3000, 1.0, 2.0), (40000, 1.0, 2.0), (40000, 2.0, 2.0),
                                           (20000, 0.5, 0.4)])
def test_grid_forces_vs_oracle(orc, n, cell, cutoff):
    from nbody_b200 import ics
    st = ics.uniform_box(n, -8, 8, seed=23)
    d, calc = _run(st, cell, cutoff, eps=0.05, G=2.0)
    out = d.numpy()
    ref_grid = orc.grid_build(st["pos_x"], st["pos_y"], st["pos_z"], cell)
    ax, ay, az, pairs = orc.spatial_forces(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], ref_grid, cutoff, 2.0,
                                           0.05)
    mag = np.sqrt(ax.astype(np.float64) ** 2 + ay ** 2 + az ** 2)
    err = np.sqrt((out["acc_x"] - ax.astype(np.float64)) ** 2 + (out["acc_y"] - ay) ** 2 + (out["acc_z"] - az) ** 2)
    # 1e-5 relative; bodies whose net short-range pull nearly cancels are held to 1e-5 of the RMS
    # acceleration instead (their own |a| is far below the sum of the |terms| that rounding acts on)
    rms = float(np.sqrt(np.mean(mag ** 2)))
    ratio = err / np.maximum(np.maximum(mag, rms), 1e-30)
    assert ratio.max() <= 1e-5, f"max err ratio {ratio.max():.3e} (plain relative {np.max(err / np.maximum(mag, 1e-30)):.3e})"
    assert np.isfinite(out["acc_x"]).all()
    # in-cutoff pairs, counted on demand by nb_sh_stats with the force sweep's own test: exact (the oracle evaluates
    # r^2 with the contraction of the reference's binary, as the kernel does), and stable when asked twice
    import ctypes as C
    import nbody_b200 as nb
    from nbody_b200 import _lib
    got = C.c_uint64(0)
    for _ in range(2):
        _lib.check(_lib.lib().nb_sh_stats(nb.get_context().handle, C.byref(got)))
        assert got.value == pairs


def test_pair_count_follows_the_last_force_evaluation(orc):
    """nb_sh_stats counts on demand: 0 on a grid no forces were evaluated on, the count of the LAST evaluation's
    cutoff afterwards, and a rebuild forgets it."""
    import ctypes as C
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    L, ctx = _lib.lib(), nb.get_context()
    st = ics.uniform_box(5000, -4, 4, seed=3)
    h = nb.ParticleDataManager.fromState(st)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    got = C.c_uint64(123)
    _lib.check(L.nb_sh_build(ctx.handle, C.byref(d.c), 1.0, None))
    _lib.check(L.nb_sh_stats(ctx.handle, C.byref(got)))
    assert got.value == 0
    grid = orc.grid_build(st["pos_x"], st["pos_y"], st["pos_z"], 1.0)
    for cutoff in (2.0, 0.7, 0.0):
        _lib.check(L.nb_sh_forces(ctx.handle, C.byref(d.c), cutoff, 1.0, 0.05, None))
        _lib.check(L.nb_sh_stats(ctx.handle, C.byref(got)))
        want = orc.spatial_forces(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], grid, cutoff, 1.0, 0.05)[3]
        assert got.value == want, cutoff
    _lib.check(L.nb_sh_build(ctx.handle, C.byref(d.c), 1.0, None))
    _lib.check(L.nb_sh_stats(ctx.handle, C.byref(got)))
    assert got.value == 0


def test_cell_index_anchor(orc):
    """tests/test_spatial_hash.cpp:38-51"""
    assert orc.cell_index(0.5, 0.5, 0.5, 2.0) == (0, 0, 0)
    assert orc.cell_index(2.5, 4.5, 6.5, 2.0) == (1, 2, 3)


def test_grid_too_large_raises():
    """src/cuda/force_spatial_hash.cu:252-254: std::runtime_error above 1e8 cells."""
    from nbody_b200 import ics
    st = ics.uniform_box(100, -300, 300, seed=1)
    with pytest.raises(RuntimeError, match="too large"):
        _run(st, 1.0, 2.0)
