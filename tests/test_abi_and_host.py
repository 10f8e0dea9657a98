"""CPU: the C-ABI library loads and exports every symbol include/nbody_b200.h declares; host-side
logic (validation, sharding layout, ICs) behaves like the reference's."""
import ctypes as C
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    txt = open(os.path.join(ROOT, "include", "nbody_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(nb_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    from nbody_b200 import _lib
    L = C.CDLL(_lib.LIB_PATH)
    names = _declared()
    assert len(names) >= 35
    for nm in names:
        assert hasattr(L, nm), f"{nm} declared in include/nbody_b200.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), set(names) ^ set(_lib.SYMBOLS)
    assert _lib.lib().nb_version() == 100


def test_nb_particles_layout_matches_reference_particledata():
    """include/nbody/types.hpp:234-276: 13 pointers then count; 112 bytes, count at 104."""
    from nbody_b200._lib import NbParticles
    assert C.sizeof(NbParticles) == 112 and NbParticles.count.offset == 104 and NbParticles.mass.offset == 96
    assert NbParticles.pos_x.offset == 0 and NbParticles.acc_old_x.offset == 72


def test_no_silent_fallback_without_gpu():
    """On a box without a GPU the product must fail loudly (no CPU path)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from nbody_b200 import _lib
    h = C.c_void_p()
    rc = _lib.lib().nb_context_create(0, 0, C.byref(h))
    assert rc == _lib.NB_ERR_NO_DEVICE and not h.value
    assert "no CPU fallback" in _lib.last_error()
    with pytest.raises(_lib.NoDeviceError):
        _lib.check(rc)
    # null-context calls are rejected, not crashed
    assert _lib.lib().nb_direct_forces(None, None, 1.0, 1e-4, None) == _lib.NB_ERR_INVALID


def test_product_does_not_import_oracle():
    """Only tests/, __graft_entry__.smoke() and bench.py's CPU legs may touch oracle/: the product
    (nbody_b200/, include/) never includes, imports, links or dlopens it."""
    bad = re.compile(r'#\s*include\s*[<"][^>"]*oracle|^\s*(from|import)\s+oracle|liboracle|libnbody_ref|dlopen', re.M)
    for top in ("nbody_b200", "include", "examples"):
        for dp, _, files in os.walk(os.path.join(ROOT, top)):
            for fn in files:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")) or fn == "Makefile":
                    txt = open(os.path.join(dp, fn), errors="replace").read()
                    assert not bad.search(txt), f"{os.path.join(dp, fn)} reaches into the oracle"


# ---- validation: tests/test_validation.cpp / src/utils/error_handling.cpp:46-123 -------------------
def test_config_validation_like_reference():
    import nbody_b200 as nb
    from nbody_b200.api import validateSimulationConfig as v
    v(nb.SimulationConfig())
    bad = [dict(particle_count=0), dict(particle_count=100000001), dict(dt=0.0), dict(dt=-1.0), dict(dt=1.5),
           dict(dt=float("nan")), dict(softening=-0.1), dict(softening=float("inf")), dict(G=0.0), dict(G=float("nan")),
           dict(cuda_block_size=0), dict(cuda_block_size=1025),
           dict(force_method=nb.ForceMethod.BARNES_HUT, barnes_hut_theta=2.5),
           dict(force_method=nb.ForceMethod.BARNES_HUT, barnes_hut_theta=-0.1),
           dict(force_method=nb.ForceMethod.SPATIAL_HASH, spatial_hash_cell_size=0.0),
           dict(force_method=nb.ForceMethod.SPATIAL_HASH, spatial_hash_cutoff=-1.0)]
    for kw in bad:
        with pytest.raises(nb.ValidationException, match="Validation Error"):
            v(nb.SimulationConfig(**kw))
    # boundaries that are legal in the reference
    v(nb.SimulationConfig(dt=1.0, softening=0.0, cuda_block_size=1024, particle_count=100000000))
    v(nb.SimulationConfig(force_method=nb.ForceMethod.BARNES_HUT, barnes_hut_theta=2.0))
    # theta is only checked for Barnes-Hut (error_handling.cpp:51-53)
    v(nb.SimulationConfig(barnes_hut_theta=5.0))


def test_factory_defaults_like_reference():
    """src/cuda/force_spatial_hash.cu:380-401: the bare factory leaves Direct at G = 1, eps = 0.01."""
    import nbody_b200 as nb
    cfg = nb.SimulationConfig(G=3.0, softening=0.5, barnes_hut_theta=0.7, spatial_hash_cell_size=2.0,
                              spatial_hash_cutoff=3.0)
    d = nb.createForceCalculator(nb.ForceMethod.DIRECT_N2, cfg)
    assert d.getMethod() == nb.ForceMethod.DIRECT_N2 and d.getGravitationalConstant() == 1.0
    assert abs(d.getSofteningParameter() - 0.01) < 1e-9
    b = nb.createForceCalculator(nb.ForceMethod.BARNES_HUT, cfg)
    assert b.getGravitationalConstant() == 3.0 and b.getSofteningParameter() == 0.5 and abs(b.getTheta() - 0.7) < 1e-7
    s = nb.createForceCalculator(nb.ForceMethod.SPATIAL_HASH, cfg)
    assert s.getCellSize() == 2.0 and s.getCutoffRadius() == 3.0 and s.getGravitationalConstant() == 3.0
    d.setSofteningParameter(0.25)
    assert float(d.softening_eps2_) == 0.0625


# ---- ICs: tests/test_particle_data.cpp:40-207 (bounds only) --------------------------------------------
def test_ic_bounds_like_reference():
    from nbody_b200 import ics
    s = ics.uniform_box(5000, -10, 10)
    for k in ("pos_x", "pos_y", "pos_z"):
        assert s[k].min() >= -10 and s[k].max() <= 10
    assert (s["mass"] == 1).all() and not s["vel_x"].any()
    s = ics.uniform_sphere(5000, 10.0)
    r = np.sqrt(s["pos_x"] ** 2 + s["pos_y"] ** 2 + s["pos_z"] ** 2)
    assert r.max() <= 10.0 * (1 + 1e-6)
    s = ics.disk(5000, 10.0, 1.0, 0.5)
    assert np.sqrt(s["pos_x"] ** 2 + s["pos_y"] ** 2).max() <= 10.0 * (1 + 1e-6) and np.abs(s["pos_z"]).max() <= 0.5
    for st in (s, ics.plummer(1000)):
        for v in st.values():
            assert v.dtype == np.float32 and v.flags["C_CONTIGUOUS"]


def test_plummer_is_in_virial_equilibrium(orc):
    from nbody_b200 import ics
    s = ics.plummer(8192, seed=42)
    assert abs(s["mass"].sum(dtype=np.float64) - 1.0) < 1e-6
    ke = orc.kinetic(s["vel_x"], s["vel_y"], s["vel_z"], s["mass"])
    pe = orc.potential(s["pos_x"], s["pos_y"], s["pos_z"], s["mass"], 1.0, 0.0)
    assert abs(-2 * ke / pe - 1.0) < 0.05          # virial ratio
    assert abs((ke + pe) + 0.25) < 0.02            # Henon units: E = -1/4
    assert s["pos_x"].mean() < 1e-6 and ics.plummer(100, seed=1)["pos_x"][0] == ics.plummer(100, seed=1)["pos_x"][0]


def test_shard_layout_properties():
    from nbody_b200.dist import shard_layout
    for n in (1, 7, 4096, 32768, 1 << 20, (1 << 22) + 3):
        for world in (1, 2, 3, 4, 8):
            if world > n:
                continue
            ends, pads = 0, set()
            for r in range(world):
                b, c, p = shard_layout(n, r, world)
                assert b == ends and c <= p and p % (256 if n < 32768 else 1024) == 0
                ends = b + c
                pads.add(p)
            assert ends == n and len(pads) == 1
    with pytest.raises(ValueError):
        shard_layout(10, 2, 2)



@pytest.mark.parametrize("std", ["c++17", "c++20"])
def test_host_headers_compile_as_the_reference_standard_and_as_cxx20(std):
    """The reference builds as C++17 (CMakeLists.txt:25); the host layer is written for C++20 and must stay usable
    from both, warning-free."""
    import shutil
    import subprocess
    if shutil.which("g++") is None:
        pytest.skip("no g++")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    for src in ("tests/cpp/test_host_api.cpp", "examples/benchmark_main.cpp"):
        r = subprocess.run(["g++", f"-std={std}", "-O0", "-Wall", "-Wextra", "-Werror", "-fsyntax-only",
                            "-I" + os.path.join(root, "include"), os.path.join(root, src)], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr[:2000]
