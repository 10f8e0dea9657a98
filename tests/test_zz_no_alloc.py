"""GPU: nb_direct_forces_static, the entry point that allocates nothing (scratch in a file-scope static __device__
array, include/nbody_b200.h), against nb_direct_forces on the same inputs, through the C ABI.  No counterpart in the
reference.  The file sorts last on purpose;
it passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

from tests.conftest import rel_err

pytestmark = pytest.mark.gpu

TOL = 1e-5  # the direct-N^2 tolerance of tests/test_gpu_direct.py (BASELINE.json north_star)
G, EPS2 = 1.0, 1e-4
ACC = ("acc_x", "acc_y", "acc_z")


def _device_data(state):
    import nbody_b200 as nb
    h = nb.ParticleDataManager.fromState(state)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    return d


def _both(st, eps2=EPS2):
    """accelerations of the allocating and of the static entry point on the same bodies"""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib
    L, ctx = _lib.lib(), nb.get_context()
    a, b = _device_data(st), _device_data(st)
    _lib.check(L.nb_direct_forces(ctx.handle, C.byref(a.c), G, eps2, None))
    for k in ACC:  # the static call has to overwrite these
        b.arrays[k].fill_(float("nan"))
    _lib.check(L.nb_direct_forces_static(ctx.handle, C.byref(b.c), G, eps2, None))
    torch.cuda.synchronize()
    return a.numpy(), b.numpy()


def test_static_matches_allocating_entry_point_two_calls_in_a_row(orc):
    from nbody_b200 import ics
    # first a large call whose partial sums fill the array, then a smaller one that finds them there; then the
    # large one again over the small one's leftovers
    first = ics.plummer(40000, seed=7)
    second = ics.plummer(5000, seed=8)
    for st, exact in ((first, False), (second, True), (first, False)):
        ref, got = _both(st)
        assert all(np.isfinite(got[k]).all() for k in ACC)
        e = rel_err(tuple(got[k] for k in ACC), tuple(ref[k] for k in ACC))
        assert e.max() <= TOL, f"n={st['mass'].size}: max rel diff to nb_direct_forces {e.max():.3e}"
        if exact:  # same chunk plan at n <= 16384: the same bits
            for k in ACC:
                assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), k
        for k in ("pos_x", "pos_y", "pos_z", "mass", "vel_x", "vel_y", "vel_z"):
            assert np.array_equal(got[k], ref[k]), f"{k} must not be touched"
    # and against the fp64 oracle, as the allocating path is checked in test_gpu_direct.py
    o = orc.direct_f64(second["pos_x"], second["pos_y"], second["pos_z"], second["mass"], G, EPS2)
    _, got = _both(second)
    assert rel_err(tuple(got[k] for k in ACC), o).max() <= TOL


@pytest.mark.parametrize("n", [1, 2, 257, 16384, 32767, 32768, 65536])
def test_static_sizes_up_to_the_maximum(n):
    from nbody_b200 import ics
    st = ics.plummer(n, seed=11) if n > 2 else ics.uniform_box(n, -1, 1, seed=3)
    ref, got = _both(st)
    if n == 1:
        assert all(got[k][0] == 0 for k in ACC)
        return
    e = rel_err(tuple(got[k] for k in ACC), tuple(ref[k] for k in ACC))
    assert e.max() <= TOL, f"n={n}: {e.max():.3e}"
    if n <= 16384:
        for k in ACC:
            assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), k


def test_static_tiny_softening_uses_the_guarded_kernel():
    from nbody_b200 import ics
    st = ics.uniform_box(3000, -1, 1, seed=5)
    ref, got = _both(st, eps2=0.0)
    for k in ACC:
        assert np.isfinite(got[k]).all()
        assert np.array_equal(got[k].view(np.uint32), ref[k].view(np.uint32)), k


def test_static_rejects_sizes_beyond_the_maximum():
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    L, ctx = _lib.lib(), nb.get_context()
    n = 65536 + 1
    d = _device_data(ics.uniform_box(n, -1, 1, seed=1))
    for k in ACC:
        d.arrays[k].fill_(123.0)
    rc = L.nb_direct_forces_static(ctx.handle, C.byref(d.c), G, EPS2, None)
    assert rc == -1  # NB_ERR_INVALID
    msg = L.nb_last_error().decode()
    assert "65536" in msg and "maximum" in msg, msg
    torch.cuda.synchronize()
    assert all((d.numpy()[k] == 123.0).all() for k in ACC), "nothing may be launched on a rejected call"
    # and the array still works afterwards
    ref, got = _both(ics.plummer(1000, seed=2))
    assert np.array_equal(got["acc_x"].view(np.uint32), ref["acc_x"].view(np.uint32))


def test_static_call_allocates_nothing():
    """free device memory and the library's scratch accounting are the same before and after calls at the maximum
    size (after one warm call: the first launch of a kernel loads its module)"""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    L, ctx = _lib.lib(), nb.get_context()
    d = _device_data(ics.plummer(65536, seed=4))
    small = _device_data(ics.plummer(3000, seed=4))
    for x in (d, small):  # warm: both kernel configurations
        _lib.check(L.nb_direct_forces_static(ctx.handle, C.byref(x.c), G, EPS2, None))
    torch.cuda.synchronize()
    gen0, bytes0 = L.nb_scratch_generation(ctx.handle), L.nb_scratch_bytes(ctx.handle)
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        for x in (d, small):
            _lib.check(L.nb_direct_forces_static(ctx.handle, C.byref(x.c), G, EPS2, None))
    torch.cuda.synchronize()
    assert L.nb_scratch_generation(ctx.handle) == gen0 and L.nb_scratch_bytes(ctx.handle) == bytes0
    assert torch.cuda.mem_get_info()[0] == free0


def test_static_call_can_be_captured_and_replayed():
    """The use case the entry point exists for: after one call outside capture (address and occupancy look-ups), the
    call is captured on a stream -- nothing in it allocates -- and the graph replays to the same bits as the eager
    call, also after the inputs changed in place."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    L, ctx = _lib.lib(), nb.get_context()
    st = ics.plummer(20000, seed=13)
    d = _device_data(st)
    _lib.check(L.nb_direct_forces_static(ctx.handle, C.byref(d.c), G, EPS2, None))   # first call: outside capture
    torch.cuda.synchronize()
    stream, graph = C.c_void_p(), C.c_void_p()
    _lib.check(L.nb_stream_create(ctx.handle, C.byref(stream)))
    try:
        gen0 = L.nb_scratch_generation(ctx.handle)
        _lib.check(L.nb_graph_begin_capture(ctx.handle, stream))
        rc = L.nb_direct_forces_static(ctx.handle, C.byref(d.c), G, EPS2, stream)
        _lib.check(L.nb_graph_end_capture(ctx.handle, stream, C.byref(graph)))
        _lib.check(rc)
        assert L.nb_scratch_generation(ctx.handle) == gen0
        for shift in (0.0, 0.25):
            d.arrays["pos_x"].add_(shift)          # same buffers, new contents
            for k in ACC:
                d.arrays[k].fill_(float("nan"))
            torch.cuda.synchronize()               # torch's stream and the capture stream do not order each other
            _lib.check(L.nb_graph_launch(graph, stream))
            _lib.check(L.nb_synchronize(ctx.handle, stream))
            got = {k: d.arrays[k].clone() for k in ACC}
            _lib.check(L.nb_direct_forces(ctx.handle, C.byref(d.c), G, EPS2, None))
            torch.cuda.synchronize()
            for k in ACC:
                assert torch.isfinite(got[k]).all()
                # 20 000 bodies: beyond 16 384 the two entry points group the partial sums differently
                a, b = got[k].double(), d.arrays[k].double()
                assert ((a - b).abs().max() <= 1e-5 * b.abs().max()).item(), k
    finally:
        if graph:
            L.nb_graph_destroy(graph)
        L.nb_stream_destroy(ctx.handle, stream)
