"""The reference's OWN test sources for the hot path (tests/test_{force_calculation,integrator,barnes_hut,
spatial_hash,particle_data,serialization,performance_observability,validation,camera,color_mapping}.cpp, 71 cases incl. the RapidCheck
properties at 100 cases each), compiled UNMODIFIED against this repo's drop-in headers and libnbody_b200.so by
oracle/build_ref_tests.sh (GoogleTest / RapidCheck / glm stand-ins: tests/cpp/shim), run on the GPU.

One case cannot pass for any fp32 implementation, the reference's included: Integrator.EnergyConservationProperty
starts a two-body system at v = sqrt(G m / 2r), for which KE = 0.1 and PE = -0.1 cancel to E0 ~ 5e-8 (the softening
residue), and then bounds |dE| / |E0| by 1 %; fp32 rounding of the state alone moves E by ~1e-8.  The second test
below runs the reference's own kernels on exactly that case and shows the same violation."""
import os
import re
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "ref_tests_on_b200")
ILL_CONDITIONED = {"Integrator.EnergyConservationProperty"}


def test_reference_test_sources_pass_against_the_drop_in():
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/ref_tests_on_b200 not built (needs /root/reference at build time)")
    r = subprocess.run([EXE], capture_output=True, text=True, timeout=900)
    out = r.stdout
    m = re.search(r"(\d+) tests ran, (\d+) passed, (\d+) failed", out)
    assert m, out[-2000:]
    ran, passed, failed = (int(x) for x in m.groups())
    bad = set(re.findall(r"^\[  FAILED  \] (\S+)$", out, flags=re.M))
    assert ran >= 71, out[-2000:]
    assert bad <= ILL_CONDITIONED, "reference tests failing against the drop-in: %s\n%s" % (sorted(bad), out[-3000:])
    assert passed >= ran - len(ILL_CONDITIONED)


def test_ill_conditioned_reference_case_also_fails_on_the_reference_kernels():
    """tests/test_integrator.cpp:90-162 evaluated for both implementations: E0 ~ 5e-8, so the 1 % relative bound asks
    for |dE| < 5e-10, below fp32 resolution of the state; what matters -- |dE| against the energy scale KE = 0.1 --
    is tiny for both."""
    import nbody_b200 as nb
    from nbody_b200 import ics
    from oracle import ref_bindings as ref
    st = ics.empty_state(2)
    r, G, m = np.float32(5.0), 1.0, 1.0
    v = np.float32(np.sqrt(np.float32(G * m) / (2 * r)))
    st["pos_x"][:] = [-r, r]
    st["vel_y"][:] = [-v, v]
    st["mass"][:] = m
    h = nb.ParticleDataManager.fromState(st)
    d = nb.ParticleData()
    nb.ParticleDataManager.allocateDevice(d, 2)
    nb.ParticleDataManager.copyToDevice(d, h)
    calc = nb.DirectForceCalculator()
    calc.setSofteningParameter(0.01)
    calc.computeForces(d)
    integ = nb.Integrator()
    e0 = integ.computeTotalEnergy(d, G, 0.01)
    for _ in range(100):
        integ.integrate(d, calc, 1e-3)
    e1 = integ.computeTotalEnergy(d, G, 0.01)
    assert abs(e0) < 1e-7                      # the test's denominator is a rounding residue
    assert abs(e1 - e0) < 1e-6 * 0.1           # energy is conserved to 1e-6 of the kinetic energy
    if ref.available():
        rs = {k: a.copy() for k, a in st.items()}
        _, en = ref.run(rs, 0, 1e-3, G, 0.01, steps=100, initial_forces=True, want_energy=True)
        r0, r1 = np.float32(en[0]) + np.float32(en[1]), np.float32(en[2]) + np.float32(en[3])
        assert abs(r0) < 2e-7
        assert abs(float(r1) - float(r0)) < 1e-6 * 0.1
        # its own bound, |dE| / |E0| < 0.01, does not hold for the reference either (or E0 rounds to exactly 0)
        assert r0 == 0 or abs(float(r1) - float(r0)) / abs(float(r0)) >= 0.01
