"""GPU parity: the CUDA splat renderer (nb_render_splats through the Renderer mirror) against the CPU
oracle on the same inputs: the RGBA8 framebuffer and the fragment count are bit-exact."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_data(state):
    import nbody_b200 as nb
    h = nb.ParticleDataManager.fromState(state)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, h.count)
    nb.ParticleDataManager.copyToDevice(d, h)
    return d


def _render_both(orc, st, w, h, mode, point_size, cam_pos=(0, 0, 50)):
    import torch
    import nbody_b200 as nb
    r = nb.Renderer()
    r.initialize(w, h)
    r.getCamera().setPosition(cam_pos)
    r.setColorMode(mode)
    r.setPointSize(point_size)
    d = _device_data(st)
    r.render(d)
    torch.cuda.synchronize()
    got = r.readPixels()
    cam = r.getCamera()
    want, frags = orc.render_splats(st, cam.getViewMatrix(), cam.getProjectionMatrix(), point_size, 100.0, 10.0, mode,
                                    w, h)
    return r, got, want, frags


@pytest.mark.parametrize("mode", [0, 1, 2])
@pytest.mark.parametrize("n,w,h,point_size", [(4096, 1280, 720, 2.0), (50000, 320, 180, 2.0), (3000, 256, 256, 400.0),
                                              (1, 64, 36, 2.0), (200000, 160, 90, 60.0)])
def test_framebuffer_bit_exact(orc, n, w, h, mode, point_size):
    from nbody_b200 import ics
    st = ics.plummer(n, seed=5) if n > 10000 else ics.uniform_sphere(n, 10.0, 1.0, seed=5)
    if n > 10000:   # spread the Plummer sphere over the view, give it visible speeds
        for k in ("pos_x", "pos_y", "pos_z"):
            st[k] = (st[k] * np.float32(8.0)).astype(np.float32)
        for k in ("vel_x", "vel_y", "vel_z"):
            st[k] = (st[k] * np.float32(6.0)).astype(np.float32)
    r, got, want, frags = _render_both(orc, st, w, h, mode, point_size)
    assert r.fragments() == frags
    assert np.array_equal(got, want), f"{(got != want).any(axis=2).sum()} pixels differ"
    assert (got != got[0, 0]).any() or n == 1


def test_camera_inside_the_cloud_and_empty(orc, tmp_path):
    """Bodies behind the camera and outside the frustum are culled; heavy overdraw near the eye
    (sprites clamped to the 63-px maximum); zero bodies = cleared framebuffer; PPM output."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    st = ics.uniform_sphere(20000, 10.0, 1.0, seed=9)
    r, got, want, frags = _render_both(orc, st, 200, 120, 2, 30.0, cam_pos=(0.5, 0.25, 3.0))
    assert r.fragments() == frags and frags > 20000
    assert np.array_equal(got, want)
    out = tmp_path / "frame.ppm"
    r.savePPM(str(out))
    raw = out.read_bytes()
    assert raw.startswith(b"P6\n200 120\n255\n") and len(raw) == len(b"P6\n200 120\n255\n") + 200 * 120 * 3
    empty = {k: np.zeros(0, np.float32) for k in st}
    r2, got2, want2, frags2 = _render_both(orc, empty, 64, 36, 0, 2.0)
    assert frags2 == 0 and np.array_equal(got2, want2) and (got2 == np.array([5, 5, 13, 255], np.uint8)).all()


def test_interop_pack_matches_interleaving():
    """K16: SoA -> interleaved xyz (src/render/cuda_gl_interop.cpp:7-28)."""
    import torch
    import nbody_b200 as nb
    from nbody_b200 import _lib, ics
    st = ics.uniform_sphere(10007, 10.0, 1.0, seed=3)
    st["vel_x"] = np.arange(10007, dtype=np.float32)
    d = _device_data(st)
    pos = torch.empty(3 * 10007, dtype=torch.float32, device="cuda")
    vel = torch.empty(3 * 10007, dtype=torch.float32, device="cuda")
    _lib.check(_lib.lib().nb_interop_pack(nb.get_context().handle, C.byref(d.c), C.c_void_p(pos.data_ptr()),
                                          C.c_void_p(vel.data_ptr()), None))
    torch.cuda.synchronize()
    assert np.array_equal(pos.cpu().numpy().reshape(-1, 3), np.stack([st["pos_x"], st["pos_y"], st["pos_z"]], 1))
    assert np.array_equal(vel.cpu().numpy().reshape(-1, 3), np.stack([st["vel_x"], st["vel_y"], st["vel_z"]], 1))


def test_particle_system_interop_buffers_follow_the_state():
    """CudaGLInterop headless (src/render/cuda_gl_interop.cpp, particle_system.cpp:96-99, 123-126, 320-345): the
    interleaved xyz buffers are refreshed by update() and follow a re-initialisation."""
    import nbody_b200 as nb
    sys_ = nb.ParticleSystem()
    sys_.initializeWithDistribution(3000, nb.InitDistribution.DISK)
    assert sys_.getInterop() is None and sys_.getParticleCount() == 3000
    sys_.initializeInterop()
    io = sys_.getInterop()
    assert io.getParticleCount() == 3000 and io.getPositionVBO() == 0 and not io.isMapped()

    def xyz(prefix):
        a = sys_.getState().arrays
        return np.stack([a[prefix + "_x"], a[prefix + "_y"], a[prefix + "_z"]], axis=1).reshape(-1)

    before = xyz("pos")
    assert io.verifyDataIntegrity(before)
    for _ in range(3):
        sys_.update(1e-3)
    assert not io.verifyDataIntegrity(before)
    assert io.verifyDataIntegrity(xyz("pos"))
    vel = io.mapVelocityBuffer()
    assert io.isVelocityMapped() and np.array_equal(vel.cpu().numpy(), xyz("vel"))
    io.unmapAllBuffers()
    assert not io.isVelocityMapped()
    sys_.initializeWithDistribution(1000, nb.InitDistribution.UNIFORM)
    assert sys_.getInterop().getParticleCount() == 1000 and sys_.getInterop().verifyDataIntegrity(xyz("pos"))
    assert sys_.getForceMethod() == nb.ForceMethod.DIRECT_N2 and sys_.getTimeStep() == nb.SimulationConfig().dt
