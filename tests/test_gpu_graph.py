"""GPU: CUDA-graph replay of the steady-state step leaves exactly the
state the eager step leaves, for the fused direct step and for the Barnes-Hut step."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _system(method, n, graph):
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.plummer(n, seed=11)
    s = nb.ParticleSystem()
    s.config_ = nb.SimulationConfig(particle_count=n, softening=0.01, dt=1e-3, force_method=method)
    s.setState(nb.SimulationState(n, 0.0, 1e-3, 1.0, 0.01, method,
                                  {k: st[k] for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "mass")}))
    s.setUseCudaGraph(graph)
    return s


@pytest.mark.parametrize("method_name,n", [("DIRECT_N2", 4096), ("DIRECT_N2", 50000), ("BARNES_HUT", 30000)])
def test_graph_replay_is_bit_identical_to_eager(method_name, n):
    import torch
    import nbody_b200 as nb
    method = getattr(nb.ForceMethod, method_name)
    eager, graphed = _system(method, n, False), _system(method, n, True)
    for _ in range(6):       # step 1 eager, step 2 captured and launched, steps 3.. replayed
        eager.update(1e-3)
        graphed.update(1e-3)
    torch.cuda.synchronize()
    assert graphed.graphNodeCount() >= 3
    a, b = eager.getDeviceData().numpy(), graphed.getDeviceData().numpy()
    for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z", "acc_old_x"):
        assert np.array_equal(a[k], b[k]), k
    # a changed parameter drops the graph: one eager step, a new capture, still identical
    eager.update(5e-4)
    graphed.update(5e-4)
    eager.update(5e-4)
    graphed.update(5e-4)
    torch.cuda.synchronize()
    a, b = eager.getDeviceData().numpy(), graphed.getDeviceData().numpy()
    assert np.array_equal(a["pos_x"], b["pos_x"]) and np.array_equal(a["vel_z"], b["vel_z"])
    assert abs(graphed.getSimulationTime() - eager.getSimulationTime()) < 1e-12


def test_capture_on_the_default_stream_is_refused():
    import nbody_b200 as nb
    from nbody_b200 import _lib
    with pytest.raises(nb.ValidationException, match="non-default stream"):
        _lib.check(_lib.lib().nb_graph_begin_capture(nb.get_context().handle, None))


def test_graph_is_dropped_when_another_user_reallocates_the_context_scratch():
    """A captured step holds raw pointers into context-owned scratch.  The canonical loop -- Barnes-Hut with
    setUseCudaGraph(true) plus Renderer::render each frame -- shares the sort scratch: a frame with more fragments
    than any sort before reallocates it, and a replay would then read and write freed memory.  The library counts
    reallocations (nb_scratch_generation); a stale graph is refused (NB_ERR_STALE_GRAPH) and the host layer drops it,
    runs one eager step and captures again.  Interleaving renders of growing size must leave exactly the eager state."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib
    n = 20000
    method = nb.ForceMethod.BARNES_HUT
    eager, graphed = _system(method, n, False), _system(method, n, True)
    L, ctx = _lib.lib(), nb.get_context()
    r = nb.Renderer()
    r.initialize(640, 360)
    r.getCamera().setPosition((0, 0, 6))
    gens, stale_seen = [], 0
    for step, point_size in enumerate([2.0, 2.0, 2.0, 40.0, 40.0, 40.0, 400.0, 400.0, 400.0]):
        eager.update(1e-3)
        if graphed.graph_ and L.nb_graph_is_stale(graphed.graph_):
            stale_seen += 1
        graphed.update(1e-3)
        r.setPointSize(point_size)    # more and more fragments: the pixel sort outgrows the scratch sized so far
        r.render(graphed.getDeviceData())
        gens.append(int(L.nb_scratch_generation(ctx.handle)))
    torch.cuda.synchronize()
    assert gens[-1] > gens[0], "the renders were meant to reallocate the sort scratch"
    assert stale_seen >= 1, "no replay was ever attempted on a stale graph: the test does not exercise the check"
    a, b = eager.getDeviceData().numpy(), graphed.getDeviceData().numpy()
    for k in ("pos_x", "pos_y", "pos_z", "vel_x", "vel_y", "vel_z", "acc_x", "acc_y", "acc_z"):
        assert np.array_equal(a[k], b[k]), k
    # the C ABI refuses a stale graph by itself
    stream = torch.cuda.Stream()
    with torch.cuda.stream(stream):
        s = __import__("ctypes").c_void_p(stream.cuda_stream)
        _lib.check(L.nb_graph_begin_capture(ctx.handle, s))
        graphed.integrator_.integrate(graphed.getDeviceData(), graphed.force_calculator_, 1e-3)
        g = __import__("ctypes").c_void_p()
        _lib.check(L.nb_graph_end_capture(ctx.handle, s, __import__("ctypes").byref(g)))
    assert L.nb_graph_is_stale(g) == 0
    m = 2 * max(int(r.fragments()), 1 << 22)   # more pairs than any sort on this context so far: the scratch must grow
    big = torch.zeros(m, dtype=torch.int32, device="cuda")
    keys = torch.randint(0, 1 << 20, (m,), device="cuda", dtype=torch.int32)
    _lib.check(L.nb_sort_pairs(ctx.handle, keys.data_ptr(), big.data_ptr(), m, 20, None))
    torch.cuda.synchronize()
    assert L.nb_graph_is_stale(g) == 1
    assert L.nb_graph_launch(g, None) == _lib.NB_ERR_STALE_GRAPH
    _lib.check(L.nb_graph_destroy(g))
