"""GPU parity: bbox, radix sort, Morton keys, LBVH topology, centres of mass and Barnes-Hut forces
through the C ABI vs the CPU oracle (integer work bit-exact)."""
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


@pytest.mark.parametrize("n,bits", [(1, 32), (2, 32), (33, 8), (4096, 32), (4097, 30), (8191, 32), (8192, 30), (8193, 13),
                                    (16385, 1), (100003, 30), (1 << 20, 21), (3000017, 32),
                                    # every digit plan of the pass kernel: 1 x 9/10/11 bits, 2 x (6..11), 3 x (8..11)
                                    (50000, 9), (50000, 10), (70001, 11), (70001, 12), (123457, 16), (123457, 20),
                                    (123457, 22), (200003, 23), (200003, 24), (200003, 31), (20000, 2), (20000, 7)])
def test_radix_sort_stable_bit_exact(orc, n, bits):
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib
    rng = np.random.default_rng(n)
    hi = (1 << bits) - 1
    # many duplicates so stability is exercised
    keys = rng.integers(0, min(hi, max(n // 3, 1)) + 1, n, dtype=np.uint64).astype(np.uint32)
    if bits == 32 and n > 100:
        keys[::7] = rng.integers(0, 1 << 32, keys[::7].size, dtype=np.uint64).astype(np.uint32)
    if n == 8192:
        keys[:] = 0x2AAAAAAA  # one digit bin per pass holds the whole tile
    vals = np.arange(n, dtype=np.int32)
    dk = torch.from_numpy(keys.view(np.int32)).cuda()
    dv = torch.from_numpy(vals).cuda()
    ctx = nb.get_context()
    _lib.check(_lib.lib().nb_sort_pairs(ctx.handle, dk.data_ptr(), dv.data_ptr(), n, bits, None))
    torch.cuda.synchronize()
    rk, rv = orc.sort_pairs(keys, vals)
    assert np.array_equal(dk.cpu().numpy().view(np.uint32), rk)
    assert np.array_equal(dv.cpu().numpy(), rv)


def test_bbox_exact(orc):
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    for n, gen in [(1, ics.uniform_box), (1000, ics.plummer), (300001, ics.uniform_sphere)]:
        st = gen(n, seed=3) if gen is ics.plummer else gen(n, seed=3)
        st["pos_x"] -= 3.5  # negative and positive values
        d = _device_data(st)
        out = np.zeros(6, np.float32)
        _lib.check(_lib.lib().nb_bounding_box(nb.get_context().handle, C.byref(d.c),
                                              out.ctypes.data_as(C.c_void_p), None))
        assert np.array_equal(out, orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"]))


def _build(state, theta=0.5, eps=0.01):
    import nbody_b200 as nb
    d = _device_data(state)
    calc = nb.BarnesHutCalculator(theta)
    calc.setSofteningParameter(eps)
    calc.computeForces(d)
    return d, calc


@pytest.mark.parametrize("n,kind", [(1, "box"), (2, "box"), (3, "box"), (100, "plummer"), (5000, "sphere"),
                                    (65536, "plummer"), (200000, "dups"),
                                    # around the tree kernel's 512-leaf tiles and 4-leaf thread blocks
                                    (4, "box"), (5, "box"), (511, "box"), (512, "box"), (513, "box"), (1023, "plummer"),
                                    (1025, "plummer"), (4097, "sphere"),
                                    # every key equal (one point): the tree is decided by the index tie-break alone,
                                    # across tiles; and two points
                                    (3000, "coincident"), (2049, "two_points")])
def test_morton_sort_lbvh_bit_exact(orc, n, kind):
    from nbody_b200 import ics
    if kind == "plummer":
        st = ics.plummer(n, seed=5)
    elif kind == "sphere":
        st = ics.uniform_sphere(n, 10.0, seed=5)
    elif kind == "dups":  # heavy key duplication: a tight clump + far outliers, and exact coincidences
        st = ics.plummer(n, seed=6)
        for k in ("pos_x", "pos_y", "pos_z"):
            st[k] *= 1e-3
            st[k][:8] = [50, -50, 50, -50, 50, -50, 50, -50]
            st[k][100:200] = st[k][100]
    elif kind == "coincident":
        st = ics.uniform_box(n, -1, 1, seed=5)
        for k in ("pos_x", "pos_y", "pos_z"):
            st[k][:] = np.float32(0.25)
    elif kind == "two_points":
        st = ics.uniform_box(n, -1, 1, seed=5)
        for k, (a, b) in zip(("pos_x", "pos_y", "pos_z"), ((-1.0, 2.0), (0.5, 0.5), (3.0, -4.0))):
            st[k][0::2] = np.float32(a)
            st[k][1::2] = np.float32(b)
    else:
        st = ics.uniform_box(n, -1, 1, seed=5)
    d, calc = _build(st)
    t = calc.getTree().export()
    bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
    assert np.array_equal(t["bbox"], bb)
    keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
    sk, sp = orc.sort_pairs(keys, idx)
    assert np.array_equal(t["keys"], sk), "sorted Morton keys"
    assert np.array_equal(t["perm"], sp), "sort permutation"
    if n > 1:
        ref = orc.lbvh_build(sk)
        assert np.array_equal(t["left"], ref["left"])
        assert np.array_equal(t["right"], ref["right"])
        assert np.array_equal(t["parent"], ref["parent"])
        assert np.array_equal(t["prefix"], ref["prefix"])
        sx, sy, sz, sm = (st[k][sp] for k in ("pos_x", "pos_y", "pos_z", "mass"))
        com = orc.lbvh_com(ref, sx, sy, sz, sm)
        assert np.array_equal(t["node_com"], com), "centres of mass (same rounding sequence)"
        # root mass = total mass (tests/test_barnes_hut.cpp:38-62, +-0.1 %)
        assert abs(t["node_com"][0, 3] - st["mass"].sum(dtype=np.float64)) <= 1e-3 * st["mass"].sum()


@pytest.fixture
def walk(request, orc, monkeypatch):
    """the traversal's two walks (NB_BH_WALK: pair records, the default, or one node per stack entry) with the
    oracle set to restate the same one"""
    monkeypatch.setenv("NB_BH_WALK", request.param)
    orc.set_walk_pairs(request.param == "pairs")
    yield request.param
    orc.set_walk_pairs(True)


@pytest.mark.parametrize("walk", ["pairs", "nodes"], indirect=True)
@pytest.mark.parametrize("group,sup", [(64, None), (32, None), (64, 64), (32, 100)])
@pytest.mark.parametrize("n,theta", [(1, 0.5), (2, 0.5), (30, 0.1), (1000, 0.5), (20000, 0.5), (20000, 0.8),
                                     (100000, 0.5)])
def test_bh_forces_vs_lbvh_oracle(orc, n, theta, group, sup, walk, monkeypatch):
    """Same tree, same traversal groups, same operation order: only rsqrt differs (<= 1e-5).
    sup = leaves of the subtrees the groups are cut from (default 8 x group; == group: the subtrees)."""
    from nbody_b200 import ics
    monkeypatch.setenv("NB_BH_GROUP", str(group))
    if sup is None:
        monkeypatch.delenv("NB_BH_SUPER", raising=False)
    else:
        monkeypatch.setenv("NB_BH_SUPER", str(sup))
    st = ics.plummer(n, seed=8) if n > 3 else ics.uniform_box(n, -1, 1, seed=8)
    d, calc = _build(st, theta)
    out = d.numpy()
    t = calc.getTree().export()
    sp = t["perm"]
    sx, sy, sz = (st[k][sp] for k in ("pos_x", "pos_y", "pos_z"))
    tree = orc.lbvh_build(t["keys"]) if n > 1 else dict(left=t["left"], right=t["right"], prefix=t["prefix"],
                                                        first=np.zeros(1, np.int32), last=np.zeros(1, np.int32))
    ax, ay, az, inter = orc.lbvh_forces(tree, t["node_com"], sx, sy, sz, t["bbox"], theta, 1.0, 0.01, group, sup)
    got = (out["acc_x"][sp], out["acc_y"][sp], out["acc_z"][sp])
    if n == 1:
        assert got[0][0] == 0
        return
    assert rel_err(got, (ax, ay, az)).max() <= 1e-5
    assert calc.getTree().interactions() == inter


@pytest.mark.parametrize("walk", ["pairs", "nodes"], indirect=True)
@pytest.mark.parametrize("kind", ["zero_mass_runs", "coincident", "two_points"])
def test_bh_forces_degenerate_inputs_vs_lbvh_oracle(orc, kind, walk):
    """Massless bodies -- singly and in runs long enough that whole subtrees have total mass 0 and are skipped
    (src/cuda/force_barnes_hut.cu:157-158) -- and trees decided by the index tie-break alone (every key equal, two
    distinct points), for both walks: interaction counts exact, forces <= 1e-5 against the step-for-step oracle."""
    from nbody_b200 import ics
    n = 6000
    st = ics.plummer(n, seed=21)
    if kind == "zero_mass_runs":
        st["mass"][::3] = 0.0
        order = np.argsort(st["pos_x"], kind="stable")
        st["mass"][order[:700]] = 0.0   # a spatially compact massless region: whole subtrees with mass 0
    elif kind == "coincident":
        for k in ("pos_x", "pos_y", "pos_z"):
            st[k][:] = np.float32(0.25)
    else:
        for k, (a, b) in zip(("pos_x", "pos_y", "pos_z"), ((0.1, 0.9), (0.2, 0.7), (-0.3, 0.4))):
            st[k][0::2] = np.float32(a)
            st[k][1::2] = np.float32(b)
    d, calc = _build(st, 0.5)
    out = d.numpy()
    t = calc.getTree().export()
    sp = t["perm"]
    sx, sy, sz = (st[k][sp] for k in ("pos_x", "pos_y", "pos_z"))
    tree = orc.lbvh_build(t["keys"])
    ax, ay, az, inter = orc.lbvh_forces(tree, t["node_com"], sx, sy, sz, t["bbox"], 0.5, 1.0, 0.01, 64, None)
    assert calc.getTree().interactions() == inter
    got = (out["acc_x"][sp], out["acc_y"][sp], out["acc_z"][sp])
    assert all(np.isfinite(g).all() for g in got)
    ref = np.sqrt(np.asarray(ax, np.float64) ** 2 + np.asarray(ay, np.float64) ** 2 + np.asarray(az, np.float64) ** 2)
    err = np.sqrt(sum((np.asarray(g, np.float64) - np.asarray(r, np.float64)) ** 2 for g, r in zip(got, (ax, ay, az))))
    # relative to the largest acceleration: bodies whose pulls cancel exactly (coincident points) have |a| = 0
    assert err.max() <= 1e-5 * max(ref.max(), 1e-30)


@pytest.mark.parametrize("walk", ["pairs", "nodes"], indirect=True)
@pytest.mark.parametrize("cap", ["100", "100000"])
def test_bh_stack_capacity_knob_is_clamped(orc, walk, cap, monkeypatch):
    """NB_BH_STACK outside the supported range (the warp stacks, list and staging of a CTA must fit 48 KiB of shared
    memory) is clamped, not passed on to a launch that would fail; the forces do not depend on the capacity."""
    from nbody_b200 import ics
    st = ics.plummer(20000, seed=12)
    d0, calc0 = _build(st, 0.5)
    a0 = d0.numpy()
    monkeypatch.setenv("NB_BH_STACK", cap)
    d1, calc1 = _build(st, 0.5)
    a1 = d1.numpy()
    for k in ("acc_x", "acc_y", "acc_z"):
        assert np.array_equal(a0[k].view(np.uint32), a1[k].view(np.uint32)), k
    import ctypes as C
    import nbody_b200 as nb
    from nbody_b200 import _lib
    deep, capacity = C.c_int(0), C.c_int(0)
    _lib.check(_lib.lib().nb_bh_max_stack(nb.get_context().handle, C.byref(deep), C.byref(capacity)))
    assert 512 <= capacity.value <= 2560 and 0 < deep.value <= capacity.value
    monkeypatch.setenv("NB_BH_STACK", "1280")
    _build(st, 0.5)   # back to the default capacity for the tests that follow (the library keeps the last setting)


def test_bh_vs_octree_oracle_and_direct(orc):
    """Stated tolerance: at equal theta the LBVH forces agree with the reference's
    intended octree algorithm within the Barnes-Hut error itself, and are no less accurate."""
    from nbody_b200 import ics
    n, theta = 20000, 0.5
    st = ics.plummer(n, seed=12)
    d, calc = _build(st, theta)
    out = d.numpy()
    got = (out["acc_x"], out["acc_y"], out["acc_z"])
    bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
    keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
    _, sp = orc.sort_pairs(keys, idx)
    nodes, dropped = orc.octree_build(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], sp, bb, as_written=False)
    assert dropped == 0
    octf = orc.octree_forces(st["pos_x"], st["pos_y"], st["pos_z"], nodes, theta, 1.0, 0.01)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4)
    e_ours = rel_err(got, truth)
    e_oct = rel_err(octf, truth)
    e_between = rel_err(got, octf)
    rms = lambda e: float(np.sqrt(np.mean(e ** 2)))  # noqa: E731
    assert rms(e_ours) <= rms(e_oct), (rms(e_ours), rms(e_oct))
    assert rms(e_between) <= 2.0 * rms(e_oct)
    assert rms(e_ours) < 5e-3


_BH_TRUTH = {}


@pytest.mark.parametrize("theta", [0.3, 0.5, 0.8])
@pytest.mark.parametrize("shape", ["plummer", "uniform_sphere"])
def test_bh_accuracy_vs_octree_oracle_over_theta(orc, shape, theta):
    """BH criterion of north_star ("agree with the reference's Barnes-Hut at equal theta within a stated tolerance"),
    over theta and two mass distributions (the table of DESIGN.md sec. 5, as a test): against the fp64 direct sum,
    the RMS relative acceleration error of the LBVH + group-box traversal must not exceed that of the reference's
    INTENDED octree algorithm (per-body test, leaf splitting: oracle/tree.c) at the same theta, and the two force
    fields must differ by no more than twice that error (the stated tolerance)."""
    from nbody_b200 import ics
    n = 30000
    st = ics.plummer(n, seed=12) if shape == "plummer" else ics.uniform_sphere(n, 1.0, 1.0 / n, seed=12)
    if shape not in _BH_TRUTH:   # truth and octree once per shape
        truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4)
        bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
        keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
        _, sp = orc.sort_pairs(keys, idx)
        nodes, dropped = orc.octree_build(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], sp, bb, as_written=False)
        assert dropped == 0
        _BH_TRUTH[shape] = (truth, nodes)
    truth, nodes = _BH_TRUTH[shape]
    d, calc = _build(st, theta)
    out = d.numpy()
    got = (out["acc_x"], out["acc_y"], out["acc_z"])
    octf = orc.octree_forces(st["pos_x"], st["pos_y"], st["pos_z"], nodes, theta, 1.0, 0.01)
    rms = lambda e: float(np.sqrt(np.mean(e ** 2)))  # noqa: E731
    e_ours, e_oct, e_between = rms(rel_err(got, truth)), rms(rel_err(octf, truth)), rms(rel_err(got, octf))
    assert e_ours <= e_oct, (shape, theta, e_ours, e_oct)
    assert e_between <= 2.0 * e_oct, (shape, theta, e_between, e_oct)


def test_bh_reference_property_tests(orc):
    """tests/test_barnes_hut.cpp:131-201 (theta monotonic) and tests/test_spatial_hash.cpp:186-249
    (theta = 0.1, N = 30: |a| within 10 % of direct)."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.uniform_sphere(30, 10.0, seed=3)
    d, _ = _build(st, 0.1, 0.1)
    out = d.numpy()
    dd = _device_data(st)
    dc = nb.DirectForceCalculator()
    dc.setSofteningParameter(0.1)
    dc.computeForces(dd)
    do = dd.numpy()
    ab = np.sqrt(out["acc_x"] ** 2 + out["acc_y"] ** 2 + out["acc_z"] ** 2)
    adir = np.sqrt(do["acc_x"] ** 2 + do["acc_y"] ** 2 + do["acc_z"] ** 2)
    assert (np.abs(ab - adir) <= 0.1 * adir + 1e-6).all()
    st = ics.uniform_sphere(500, 10.0, seed=4)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 0.01)
    errs = []
    for theta in (0.3, 0.8):
        d, _ = _build(st, theta, 0.1)
        o = d.numpy()
        errs.append(sum(np.abs(o[k] - t).sum() for k, t in zip(("acc_x", "acc_y", "acc_z"), truth)))
    assert errs[0] <= errs[1] * 1.1


def test_bh_verlet_through_integrator(orc):
    """Integrator::integrate with the Barnes-Hut calculator (generic 4-call path)."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.plummer(4096, seed=2)
    d, calc = _build(st, 0.4)
    integ = nb.Integrator()
    e0 = integ.computeTotalEnergy(d, 1.0, 0.01)
    for _ in range(20):
        integ.integrate(d, calc, 1e-3)
    e1 = integ.computeTotalEnergy(d, 1.0, 0.01)
    assert abs(e1 - e0) / abs(e0) < 1e-3
