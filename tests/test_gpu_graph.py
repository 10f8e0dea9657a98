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
