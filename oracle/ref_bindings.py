"""ctypes bindings for oracle/_ref/libnbody_ref.so -- the UNMODIFIED reference CUDA sources compiled
for sm_100 by oracle/build_ref.sh.  TEST INFRASTRUCTURE ONLY (parity oracle on the GPU box and the
"reference kernels on the same B200" timing); needs a GPU to call anything but the host helpers.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from .bindings import OCTNODE_DTYPE

_HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.path.join(_HERE, "_ref", "libnbody_ref.so")
_L = None
_NAMES = ["pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z",
          "acc_old_x", "acc_old_y", "acc_old_z", "mass"]


def available() -> bool:
    return os.path.exists(SO)


def lib():
    global _L
    if _L is None:
        _L = C.CDLL(SO)
        _L.ref_last_error.restype = C.c_char_p
        _L.ref_morton_code.restype = C.c_uint32
        _L.ref_morton_code.argtypes = [C.c_float] * 3
    return _L


def _check(rc):
    if rc != 0:
        raise RuntimeError("reference: " + lib().ref_last_error().decode())


def _arrays(state):
    arr = (C.c_void_p * 13)()
    for k, nm in enumerate(_NAMES):
        a = state[nm]
        assert a.dtype == np.float32 and a.flags["C_CONTIGUOUS"]
        arr[k] = a.ctypes.data
    return arr


def forces(state, method=0, G=1.0, eps=0.01, theta=0.5, cell=1.0, cutoff=2.0, iters=0):
    """Runs the reference calculator; writes acc_* of `state` in place; returns mean ms if iters > 0."""
    ms = C.c_float(0)
    _check(lib().ref_forces(_arrays(state), C.c_size_t(state["pos_x"].size), int(method), C.c_float(G),
                            C.c_float(eps), C.c_float(theta), C.c_float(cell), C.c_float(cutoff), int(iters),
                            C.byref(ms)))
    return ms.value


def run(state, method=0, dt=1e-3, G=1.0, eps=0.01, theta=0.5, cell=1.0, cutoff=2.0, steps=1, initial_forces=True,
        want_energy=False):
    ms = C.c_float(0)
    en = (C.c_float * 4)()
    _check(lib().ref_run(_arrays(state), C.c_size_t(state["pos_x"].size), int(method), C.c_float(dt), C.c_float(G),
                         C.c_float(eps), C.c_float(theta), C.c_float(cell), C.c_float(cutoff), int(steps),
                         1 if initial_forces else 0, en if want_energy else None, C.byref(ms)))
    return ms.value, [float(x) for x in en]


def bh_tree(state, max_nodes=64):
    n = state["pos_x"].size
    bbox = np.zeros(6, np.float32)
    ku, ks = np.zeros(n, np.uint32), np.zeros(n, np.uint32)
    perm = np.zeros(n, np.int32)
    cnt = C.c_int(0)
    nodes = np.zeros(max_nodes, OCTNODE_DTYPE)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    _check(lib().ref_bh_tree(_arrays(state), C.c_size_t(n), p(bbox), p(ku), p(ks), p(perm), C.byref(cnt), p(nodes),
                             max_nodes))
    return dict(bbox=bbox, keys_unsorted=ku, keys=ks, perm=perm, node_count=cnt.value,
                nodes=nodes[:min(cnt.value, max_nodes)])


def sh_grid(state, cell_size):
    n = state["pos_x"].size
    bbox = np.zeros(6, np.float32)
    dims = np.zeros(3, np.int32)
    p = lambda a: a.ctypes.data_as(C.c_void_p)  # noqa: E731
    _check(lib().ref_sh_grid(_arrays(state), C.c_size_t(n), C.c_float(cell_size), p(bbox), p(dims), None, None, None,
                             None))
    total = int(dims[0]) * int(dims[1]) * int(dims[2])
    pc, si = np.zeros(n, np.int32), np.zeros(n, np.int32)
    cs, ce = np.zeros(total, np.int32), np.zeros(total, np.int32)
    _check(lib().ref_sh_grid(_arrays(state), C.c_size_t(n), C.c_float(cell_size), p(bbox), p(dims), p(pc), p(cs),
                             p(ce), p(si)))
    return dict(bbox=bbox, dims=dims, particle_cell=pc, cell_start=cs, cell_end=ce, sorted_indices=si,
                total_cells=total)


def pair_force_cpu(p1, p2, m1, m2, G, eps):
    out = np.zeros(3, np.float32)
    a, b = np.asarray(p1, np.float32), np.asarray(p2, np.float32)
    lib().ref_pair_force_cpu(a.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p), C.c_float(m1),
                             C.c_float(m2), C.c_float(G), C.c_float(eps), out.ctypes.data_as(C.c_void_p))
    return out


def morton_code(x, y, z):
    return int(lib().ref_morton_code(x, y, z))


def cell_index(x, y, z, cell_size):
    out = np.zeros(3, np.int32)
    lib().ref_cell_index(C.c_float(x), C.c_float(y), C.c_float(z), C.c_float(cell_size),
                         out.ctypes.data_as(C.c_void_p))
    return tuple(int(v) for v in out)
