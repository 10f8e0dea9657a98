"""ctypes bindings for the CPU oracle (oracle/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs.  The product package (nbody_b200) never imports it.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
f64p = np.ctypeslib.ndpointer(dtype=np.float64, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")
u32p = np.ctypeslib.ndpointer(dtype=np.uint32, flags="C_CONTIGUOUS")


class OctNode(C.Structure):
    """include/nbody/barnes_hut_tree.hpp:9-30 (76 bytes)."""

    _fields_ = [
        ("center", C.c_float * 3),
        ("half_size", C.c_float),
        ("com", C.c_float * 3),
        ("total_mass", C.c_float),
        ("children", C.c_int32 * 8),
        ("particle_index", C.c_int32),
        ("is_leaf", C.c_uint8),
        ("pad_", C.c_uint8 * 3),
        ("particle_count", C.c_int32),
    ]


OCTNODE_DTYPE = np.dtype(
    [
        ("center", np.float32, 3),
        ("half_size", np.float32),
        ("com", np.float32, 3),
        ("total_mass", np.float32),
        ("children", np.int32, 8),
        ("particle_index", np.int32),
        ("is_leaf", np.uint8),
        ("pad_", np.uint8, 3),
        ("particle_count", np.int32),
    ]
)
assert OCTNODE_DTYPE.itemsize == 76 and C.sizeof(OctNode) == 76


def build(force: bool = False) -> str:
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, s) for s in ("direct.c", "tree.c", "lbvh.c", "grid.c", "oracle.h")]
    stale = (not os.path.exists(so)) or any(
        os.path.exists(s) and os.path.getmtime(s) > os.path.getmtime(so) for s in srcs
    )
    if force or stale:
        subprocess.run(["make", "-C", _HERE, "-B", "liboracle.so"], check=True, capture_output=True)
    return so


def lib():
    global _LIB
    if _LIB is not None:
        return _LIB
    L = C.CDLL(build())
    i, f, d, vp = C.c_int, C.c_float, C.c_double, C.c_void_p
    L.orc_direct_f32.argtypes = [i, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f, f, i, i]
    L.orc_direct_f32_eps2.argtypes = L.orc_direct_f32.argtypes
    L.orc_direct_f64.argtypes = [i, f32p, f32p, f32p, f32p, f64p, f64p, f64p, d, d, vp, i]
    L.orc_pair_force.argtypes = [f32p, f32p, f, f, f, f, f32p]
    L.orc_store_acc.argtypes = [i] + [f32p] * 6
    L.orc_drift.argtypes = [i] + [f32p] * 9 + [f]
    L.orc_kick.argtypes = [i] + [f32p] * 9 + [f]
    L.orc_step_direct.argtypes = [i, C.POINTER(C.c_void_p), f, f, f]
    L.orc_run_direct.argtypes = [i, C.POINTER(C.c_void_p), f, f, f, i]
    L.orc_kinetic_f64.argtypes = [i, f32p, f32p, f32p, f32p]
    L.orc_kinetic_f64.restype = d
    L.orc_potential_f64.argtypes = [i, f32p, f32p, f32p, f32p, d, d]
    L.orc_moments_f64.argtypes = [i] + [f32p] * 7 + [C.c_void_p]
    L.orc_moments_f64.restype = None
    L.orc_potential_f64.restype = d
    L.orc_bbox.argtypes = [i, f32p, f32p, f32p, f32p]
    L.orc_expand_bits.argtypes = [C.c_uint32]
    L.orc_expand_bits.restype = C.c_uint32
    L.orc_morton_code.argtypes = [f, f, f]
    L.orc_morton_code.restype = C.c_uint32
    L.orc_morton_keys.argtypes = [i, f32p, f32p, f32p, f32p, u32p, i32p]
    L.orc_sort_pairs.argtypes = [i, u32p, i32p]
    L.orc_octree_build.argtypes = [i, f32p, f32p, f32p, f32p, i32p, f32p, i, vp, i, C.POINTER(i)]
    L.orc_octree_build.restype = i
    L.orc_octree_forces.argtypes = [i, f32p, f32p, f32p, vp, i, f, f, f, f32p, f32p, f32p]
    L.orc_lbvh_build.argtypes = [i, u32p, i32p, i32p, i32p, i32p, vp, vp]
    L.orc_lbvh_com.argtypes = [i, i32p, i32p, f32p, f32p, f32p, f32p, f32p]
    L.orc_lbvh_cell_size.argtypes = [i, f32p]
    L.orc_lbvh_cell_size.restype = f
    L.orc_lbvh_forces.argtypes = [i, i32p, i32p, i32p, i32p, i32p, f32p, f32p, f32p, f32p, f32p, f, f, f, i, i,
                                  f32p, f32p, f32p, C.POINTER(C.c_uint64)]
    L.orc_lbvh_forces_ex.argtypes = [i, i32p, i32p, i32p, i32p, i32p, f32p, f32p, f32p, f32p, f32p, f, f, f, i, i, i, f,
                                     f32p, f32p, f32p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
    L.orc_cell_index.argtypes = [f, f, f, f, i32p]
    L.orc_grid_dims.argtypes = [f32p, f, i32p]
    L.orc_assign_cells.argtypes = [i, f32p, f32p, f32p, f32p, f, i32p, i32p]
    L.orc_cell_ranges.argtypes = [i, i32p, i, i32p, i32p, i32p]
    L.orc_spatial_forces.argtypes = [i, f32p, f32p, f32p, f32p, i32p, i32p, i32p, i32p, f, f32p, f, f, f,
                                     f32p, f32p, f32p, C.POINTER(C.c_uint64)]
    L.orc_render_splats.argtypes = [i, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f32p, f, f, f, i, i, i, f, f, f32p,
                                    np.ctypeslib.ndpointer(np.uint8, flags="C_CONTIGUOUS")]
    L.orc_render_splats.restype = C.c_uint64
    L.orc_num_threads.restype = i
    L.orc_set_num_threads.argtypes = [i]
    _LIB = L
    return L


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ---- direct ----------------------------------------------------------------------------
def direct_f32(px, py, pz, m, G=1.0, eps2=1e-4, i0=0, i1=None):
    px, py, pz, m = map(_f32, (px, py, pz, m))
    n = px.size
    i1 = n if i1 is None else i1
    ax, ay, az = (np.zeros(n, np.float32) for _ in range(3))
    lib().orc_direct_f32_eps2(n, px, py, pz, m, ax, ay, az, G, eps2, i0, i1)
    return ax, ay, az


def direct_f64(px, py, pz, m, G=1.0, eps2=1e-4, idx=None):
    px, py, pz, m = map(_f32, (px, py, pz, m))
    n = px.size
    if idx is None:
        nt, ptr = n, None
    else:
        idx = np.ascontiguousarray(idx, dtype=np.int32)
        nt, ptr = idx.size, idx.ctypes.data_as(C.c_void_p)
    ax, ay, az = (np.zeros(nt, np.float64) for _ in range(3))
    lib().orc_direct_f64(n, px, py, pz, m, ax, ay, az, float(G), float(eps2), ptr, nt)
    return ax, ay, az


def pair_force(p1, p2, m1, m2, G, eps):
    out = np.zeros(3, np.float32)
    lib().orc_pair_force(_f32(p1), _f32(p2), m1, m2, G, eps, out)
    return out


def _arrays13(state):
    """state: dict with the 13 ParticleData arrays (fp32, contiguous, modified in place)."""
    names = ["pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z",
             "acc_old_x", "acc_old_y", "acc_old_z", "mass"]
    arr = (C.c_void_p * 13)()
    for k, nm in enumerate(names):
        a = state[nm]
        assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
        arr[k] = a.ctypes.data
    return arr


def run_direct(state, dt, G, eps2, steps=1):
    n = state["pos_x"].size
    lib().orc_run_direct(n, _arrays13(state), dt, G, eps2, steps)


def drift(state, dt):
    s = state
    lib().orc_drift(s["pos_x"].size, s["pos_x"], s["pos_y"], s["pos_z"], s["vel_x"], s["vel_y"], s["vel_z"],
                    s["acc_x"], s["acc_y"], s["acc_z"], dt)


def kick(state, dt):
    s = state
    lib().orc_kick(s["pos_x"].size, s["vel_x"], s["vel_y"], s["vel_z"], s["acc_old_x"], s["acc_old_y"],
                   s["acc_old_z"], s["acc_x"], s["acc_y"], s["acc_z"], dt)


def store_acc(state):
    s = state
    lib().orc_store_acc(s["pos_x"].size, s["acc_x"], s["acc_y"], s["acc_z"], s["acc_old_x"], s["acc_old_y"],
                        s["acc_old_z"])


def kinetic(vx, vy, vz, m):
    return lib().orc_kinetic_f64(vx.size, _f32(vx), _f32(vy), _f32(vz), _f32(m))


def moments(state):
    """{M, sum m x (3), sum m v (3), sum m x cross v (3), KE} in fp64."""
    out = np.zeros(11, np.float64)
    lib().orc_moments_f64(state["pos_x"].size, *(_f32(state[k]) for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y",
                                                                          "vel_z", "mass")), out.ctypes.data)
    return out


def potential(px, py, pz, m, G, eps):
    return lib().orc_potential_f64(px.size, _f32(px), _f32(py), _f32(pz), _f32(m), float(G), float(eps))


# ---- bbox / Morton / sort ----------------------------------------------------------------
def bbox(px, py, pz):
    out = np.zeros(6, np.float32)
    lib().orc_bbox(px.size, _f32(px), _f32(py), _f32(pz), out)
    return out


def morton_keys(px, py, pz, bb=None):
    px, py, pz = map(_f32, (px, py, pz))
    bb = bbox(px, py, pz) if bb is None else _f32(bb)
    keys = np.zeros(px.size, np.uint32)
    idx = np.zeros(px.size, np.int32)
    lib().orc_morton_keys(px.size, px, py, pz, bb, keys, idx)
    return keys, idx


def sort_pairs(keys, vals):
    k = np.ascontiguousarray(keys, dtype=np.uint32).copy()
    v = np.ascontiguousarray(vals, dtype=np.int32).copy()
    lib().orc_sort_pairs(k.size, k, v)
    return k, v


# ---- octree ------------------------------------------------------------------------------
def octree_build(px, py, pz, m, sorted_idx, bb, as_written=False, max_nodes=None):
    px, py, pz, m = map(_f32, (px, py, pz, m))
    n = px.size
    max_nodes = max_nodes or (8 * n + 64)
    nodes = np.zeros(max_nodes, OCTNODE_DTYPE)
    dropped = C.c_int(0)
    cnt = lib().orc_octree_build(n, px, py, pz, m, np.ascontiguousarray(sorted_idx, np.int32), _f32(bb),
                                 1 if as_written else 0, nodes.ctypes.data_as(C.c_void_p), max_nodes,
                                 C.byref(dropped))
    if cnt < 0:
        raise MemoryError("octree oracle: max_nodes too small")
    return nodes[:cnt].copy(), dropped.value


def octree_forces(px, py, pz, nodes, theta, G, eps):
    px, py, pz = map(_f32, (px, py, pz))
    n = px.size
    nodes = np.ascontiguousarray(nodes)
    ax, ay, az = (np.zeros(n, np.float32) for _ in range(3))
    lib().orc_octree_forces(n, px, py, pz, nodes.ctypes.data_as(C.c_void_p), nodes.size, theta, G, eps, ax, ay, az)
    return ax, ay, az


# ---- LBVH --------------------------------------------------------------------------------
def lbvh_build(sorted_keys):
    k = np.ascontiguousarray(sorted_keys, dtype=np.uint32)
    n = k.size
    ni = max(n - 1, 1)
    left = np.full(ni, -1, np.int32)
    right = np.full(ni, -1, np.int32)
    prefix = np.zeros(ni, np.int32)
    parent = np.full(2 * n - 1, -1, np.int32)
    first = np.zeros(ni, np.int32)
    last = np.zeros(ni, np.int32)
    lib().orc_lbvh_build(n, k, left, right, parent, prefix, first.ctypes.data_as(C.c_void_p),
                         last.ctypes.data_as(C.c_void_p))
    return dict(left=left, right=right, parent=parent, prefix=prefix, first=first, last=last)


def lbvh_com(tree, sx, sy, sz, sm):
    n = sx.size
    com = np.zeros((2 * n - 1, 4), np.float32)
    lib().orc_lbvh_com(n, tree["left"], tree["right"], _f32(sx), _f32(sy), _f32(sz), _f32(sm), com.reshape(-1))
    return com


def lbvh_cell_size(L, bb):
    return lib().orc_lbvh_cell_size(int(L), _f32(bb))


def set_walk_pairs(on):
    """1 (default): lbvh_forces restates bh_force_pairs_kernel; 0: bh_force_kernel (NB_BH_WALK=nodes)."""
    lib().orc_set_walk_pairs(1 if on else 0)


def lbvh_forces(tree, com, sx, sy, sz, bb, theta, G, eps, group=32, super_=None):
    n = sx.size
    ax, ay, az = (np.zeros(n, np.float32) for _ in range(3))
    cnt = C.c_uint64(0)
    lib().orc_lbvh_forces(n, tree["left"], tree["right"], tree["prefix"], tree["first"], tree["last"],
                          np.ascontiguousarray(com).reshape(-1),
                          _f32(sx), _f32(sy), _f32(sz), _f32(bb), theta, G, eps, group,
                          8 * group if super_ is None else super_, ax, ay, az, C.byref(cnt))
    return ax, ay, az, cnt.value


def lbvh_forces_ex(tree, com, sx, sy, sz, bb, theta, G, eps, group=64, super_=None, mode=0, alpha=0.0):
    """lbvh_forces with the grouping / acceptance variants (see oracle/lbvh.c); returns also (groups, list entries)."""
    n = sx.size
    ax, ay, az = (np.zeros(n, np.float32) for _ in range(3))
    cnt = C.c_uint64(0)
    stats = (C.c_uint64 * 2)()
    lib().orc_lbvh_forces_ex(n, tree["left"], tree["right"], tree["prefix"], tree["first"], tree["last"],
                             np.ascontiguousarray(com).reshape(-1), _f32(sx), _f32(sy), _f32(sz), _f32(bb), theta, G,
                             eps, group, 8 * group if super_ is None else super_, mode, alpha, ax, ay, az,
                             C.byref(cnt), stats)
    return ax, ay, az, cnt.value, int(stats[0]), int(stats[1])


# ---- spatial hash ------------------------------------------------------------------------
def cell_index(x, y, z, cell_size):
    out = np.zeros(3, np.int32)
    lib().orc_cell_index(x, y, z, cell_size, out)
    return tuple(int(v) for v in out)


def grid_build(px, py, pz, cell_size):
    px, py, pz = map(_f32, (px, py, pz))
    n = px.size
    bb = bbox(px, py, pz)
    dims = np.zeros(3, np.int32)
    lib().orc_grid_dims(bb, cell_size, dims)
    total = int(dims[0]) * int(dims[1]) * int(dims[2])
    pc = np.zeros(n, np.int32)
    lib().orc_assign_cells(n, px, py, pz, bb, cell_size, dims, pc)
    start = np.zeros(total, np.int32)
    end = np.zeros(total, np.int32)
    sidx = np.zeros(n, np.int32)
    lib().orc_cell_ranges(n, pc, total, start, end, sidx)
    return dict(bbox=bb, dims=dims, particle_cell=pc, cell_start=start, cell_end=end, sorted_indices=sidx,
                total_cells=total, cell_size=cell_size)


def spatial_forces(px, py, pz, m, grid, cutoff, G, eps):
    px, py, pz, m = map(_f32, (px, py, pz, m))
    n = px.size
    ax, ay, az = (np.zeros(n, np.float32) for _ in range(3))
    cnt = C.c_uint64(0)
    lib().orc_spatial_forces(n, px, py, pz, m, grid["cell_start"], grid["cell_end"], grid["sorted_indices"],
                             grid["dims"], grid["cell_size"], grid["bbox"], cutoff, G, eps, ax, ay, az,
                             C.byref(cnt))
    return ax, ay, az, cnt.value


def render_splats(state, view, proj, point_size, max_depth, max_velocity, color_mode, width, height,
                  size_min=1.0, size_max=63.0, clear=(0.02, 0.02, 0.05, 1.0)):
    """height x width x 4 uint8 (row 0 = bottom row) and the fragment count."""
    n = state["pos_x"].size
    img = np.zeros(width * height * 4, np.uint8)
    frags = lib().orc_render_splats(n, _f32(state["pos_x"]), _f32(state["pos_y"]), _f32(state["pos_z"]),
                                    _f32(state["vel_x"]), _f32(state["vel_y"]), _f32(state["vel_z"]), _f32(view),
                                    _f32(proj), point_size, max_depth, max_velocity, int(color_mode), width, height,
                                    size_min, size_max, _f32(np.asarray(clear, np.float32)), img)
    return img.reshape(height, width, 4), int(frags)


def num_threads():
    return lib().orc_num_threads()


def set_num_threads(t: int):
    lib().orc_set_num_threads(int(t))
