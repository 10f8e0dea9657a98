"""CPU: internal consistency of the oracle itself (so a bug in it cannot hide a bug in the kernels)."""
import numpy as np
import pytest

from tests.conftest import rel_err


def test_f32_oracle_matches_f64_truth(orc):
    from nbody_b200 import ics
    st = ics.plummer(3000, seed=1)
    a32 = orc.direct_f32(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4)
    a64 = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4)
    assert rel_err(a32, a64).max() < 2e-5
    idx = np.array([5, 17, 2999], np.int32)
    sub = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 1e-4, idx)
    for k in range(3):
        assert np.array_equal(sub[k], a64[k][idx])
    # momentum conservation: sum m a = 0
    for c in a64:
        assert abs((st["mass"] * c).sum()) < 1e-12


def test_morton_is_bit_interleave(orc):
    rng = np.random.default_rng(2)
    for _ in range(500):
        ix, iy, iz = (int(v) for v in rng.integers(0, 1024, 3))
        code = orc.lib().orc_morton_code((ix + 0.5) / 1024, (iy + 0.5) / 1024, (iz + 0.5) / 1024)
        want = 0
        for b in range(10):
            want |= ((ix >> b) & 1) << (3 * b + 2) | ((iy >> b) & 1) << (3 * b + 1) | ((iz >> b) & 1) << (3 * b)
        assert code == want
    assert orc.lib().orc_morton_code(1.0, 1.0, 1.0) == (1 << 30) - 1       # clamped to 1023
    assert orc.lib().orc_morton_code(-0.5, 0.0, 0.0) == 0                  # saturating conversion


def test_sort_is_stable(orc):
    rng = np.random.default_rng(3)
    keys = rng.integers(0, 50, 20000).astype(np.uint32)
    k, v = orc.sort_pairs(keys, np.arange(keys.size, dtype=np.int32))
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(v, order) and np.array_equal(k, keys[order])


def test_lbvh_invariants(orc):
    rng = np.random.default_rng(4)
    for n in (2, 3, 17, 1000):
        keys = np.sort(rng.integers(0, 1 << 30 if n > 17 else 8, n).astype(np.uint32))
        t = orc.lbvh_build(keys)
        leaf0 = n - 1
        seen = np.zeros(2 * n - 1, int)
        for i in range(n - 1):
            for c in (t["left"][i], t["right"][i]):
                seen[c] += 1
                assert t["parent"][c] == i
            lo, hi = t["first"][i], t["last"][i]
            assert lo < hi
            # children split the range
            l, r = t["left"][i], t["right"][i]
            l_hi = l - leaf0 if l >= leaf0 else t["last"][l]
            r_lo = r - leaf0 if r >= leaf0 else t["first"][r]
            assert l_hi + 1 == r_lo
            # common prefix of the range's keys
            x = int(keys[lo]) ^ int(keys[hi])
            want = 30 if x == 0 else 30 - x.bit_length()
            assert t["prefix"][i] == want
        assert seen[0] == 0 and (seen[1:] == 1).all() and t["parent"][0] == -1
        assert t["first"][0] == 0 and t["last"][0] == n - 1


def test_lbvh_com_and_forces_consistency(orc):
    from nbody_b200 import ics
    st = ics.plummer(3000, seed=6)
    bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
    keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
    sk, sp = orc.sort_pairs(keys, idx)
    t = orc.lbvh_build(sk)
    sx, sy, sz, sm = (st[k][sp] for k in ("pos_x", "pos_y", "pos_z", "mass"))
    com = orc.lbvh_com(t, sx, sy, sz, sm)
    assert abs(com[0, 3] - 1.0) < 1e-5
    np.testing.assert_allclose(com[0, :3], [(sx * sm).sum() / sm.sum(), (sy * sm).sum() / sm.sum(),
                                            (sz * sm).sum() / sm.sum()], atol=1e-5)
    truth = orc.direct_f64(sx, sy, sz, sm, 1.0, 1e-4)
    # theta -> 0 degenerates to direct summation
    a0 = orc.lbvh_forces(t, com, sx, sy, sz, bb, 0.0, 1.0, 0.01, 32)
    assert a0[3] == 3000 * 3000
    assert rel_err(a0[:3], truth).max() < 2e-5
    e3 = rel_err(orc.lbvh_forces(t, com, sx, sy, sz, bb, 0.3, 1.0, 0.01, 32)[:3], truth)
    e8 = rel_err(orc.lbvh_forces(t, com, sx, sy, sz, bb, 0.8, 1.0, 0.01, 32)[:3], truth)
    assert np.sqrt(np.mean(e3 ** 2)) < np.sqrt(np.mean(e8 ** 2)) < 0.05
    # cell size table: level 0 = largest extent, three prefix bits halve it
    s0, s3 = orc.lbvh_cell_size(0, bb), orc.lbvh_cell_size(3, bb)
    assert abs(s0 - (bb[3:] - bb[:3] + np.float32(0.001)).max()) < 1e-6 and abs(s3 - s0 / 2) < 1e-6


@pytest.mark.parametrize("n,theta,group", [(1, 0.5, 64), (2, 0.5, 64), (777, 0.5, 64), (5000, 0.5, 64),
                                           (5000, 0.8, 32), (5000, 0.0, 64)])
def test_lbvh_walks_accept_the_same_nodes(orc, n, theta, group):
    """The two walks of the traversal (child-pair records / one node per stack entry) apply the same acceptance test to
    the same nodes: identical interaction counts, sums equal to the rounding of a different order."""
    from nbody_b200 import ics
    st = ics.plummer(n, seed=9) if n > 2 else ics.uniform_box(n, -1, 1, seed=9)
    bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
    keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
    sk, sp = orc.sort_pairs(keys, idx)
    if n > 1:
        t = orc.lbvh_build(sk)
    else:
        z = np.zeros(1, np.int32)
        t = dict(left=z, right=z, prefix=z, first=z, last=z)
    sx, sy, sz, sm = (st[k][sp] for k in ("pos_x", "pos_y", "pos_z", "mass"))
    com = orc.lbvh_com(t, sx, sy, sz, sm) if n > 1 else np.array([[sx[0], sy[0], sz[0], sm[0]]], np.float32)
    try:
        orc.set_walk_pairs(True)
        a = orc.lbvh_forces(t, com, sx, sy, sz, bb, theta, 1.0, 0.01, group)
        orc.set_walk_pairs(False)
        b = orc.lbvh_forces(t, com, sx, sy, sz, bb, theta, 1.0, 0.01, group)
    finally:
        orc.set_walk_pairs(True)
    assert a[3] == b[3]
    if theta == 0.0:
        assert a[3] == n * n
    if n > 1:
        assert rel_err(a[:3], b[:3]).max() < 1e-5


def test_octree_as_written_vs_intended(orc):
    """as shipped, N = 50 -> 9 nodes / 42 bodies dropped; the intended algorithm keeps
    every body (tests/test_barnes_hut.cpp:38-62 mass conservation, :99-127 all particles in tree)."""
    from nbody_b200 import ics
    st = ics.uniform_sphere(50, 10.0, seed=42)
    bb = orc.bbox(st["pos_x"], st["pos_y"], st["pos_z"])
    keys, idx = orc.morton_keys(st["pos_x"], st["pos_y"], st["pos_z"], bb)
    _, sp = orc.sort_pairs(keys, idx)
    nodes, dropped = orc.octree_build(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], sp, bb, as_written=True)
    assert len(nodes) <= 9 and dropped == 50 - (len(nodes) - 1)
    nodes, dropped = orc.octree_build(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], sp, bb, as_written=False)
    assert dropped == 0 and abs(nodes[0]["total_mass"] - 50.0) < 0.05
    leaves = nodes[nodes["is_leaf"] == 1]
    assert sorted(leaves["particle_index"]) == list(range(50))
    f = orc.octree_forces(st["pos_x"], st["pos_y"], st["pos_z"], nodes, 0.1, 1.0, 0.1)
    truth = orc.direct_f64(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], 1.0, 0.01)
    assert rel_err(f, truth).max() < 0.1   # tests/test_spatial_hash.cpp:186-249


def test_grid_ranges_against_numpy(orc):
    from nbody_b200 import ics
    st = ics.uniform_box(5000, -10, 10, seed=9)
    g = orc.grid_build(st["pos_x"], st["pos_y"], st["pos_z"], 1.5)
    cnt = np.bincount(g["particle_cell"], minlength=g["total_cells"])
    start = np.concatenate([[0], np.cumsum(cnt)[:-1]])
    assert np.array_equal(g["cell_start"], start) and np.array_equal(g["cell_end"], start + cnt)
    assert np.array_equal(g["sorted_indices"], np.argsort(g["particle_cell"], kind="stable"))
    # cutoff larger than anything reachable by the 3x3x3 stencil == all pairs within the stencil
    ax, ay, az, pairs = orc.spatial_forces(st["pos_x"], st["pos_y"], st["pos_z"], st["mass"], g, 1e-3, 1.0, 0.1)
    assert pairs == 0 and not ax.any()


def test_moments_oracle_matches_numpy(orc):
    """orc_moments_f64 (checker of nb_diagnostics) against plain numpy fp64 sums."""
    from nbody_b200 import ics
    st = ics.plummer(5000, seed=4)
    st["pos_x"] += 0.25  # off-centre, so that the first moments are not ~0
    st["vel_y"] += 0.125
    got = orc.moments(st)
    m = st["mass"].astype(np.float64)
    x = np.stack([st[k].astype(np.float64) for k in ("pos_x", "pos_y", "pos_z")])
    v = np.stack([st[k].astype(np.float64) for k in ("vel_x", "vel_y", "vel_z")])
    want = np.concatenate([[m.sum()], (m * x).sum(1), (m * v).sum(1), (m * np.cross(x.T, v.T).T).sum(1),
                           [0.5 * (m * (v * v).sum(0)).sum()]])
    assert np.allclose(got, want, rtol=1e-12, atol=1e-15)
    assert abs(got[10] - orc.kinetic(st["vel_x"], st["vel_y"], st["vel_z"], st["mass"])) <= 1e-14
