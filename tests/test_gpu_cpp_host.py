"""GPU: the C++20 host layer (include/nbody/*.hpp) -- the reference's test cases re-expressed in C++
against the drop-in headers, and the benchmark program in the reference's JSON record shape."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


_SOURCES = {"test_host_api": "tests/cpp/test_host_api.cpp", "benchmark_main": "examples/benchmark_main.cpp"}


def _exe(name):
    """build/<name> as __graft_entry__.build() leaves it; if the tree arrived without build/, the same g++ line here
    (host code only: the CUDA library must already be there, nothing falls back)."""
    p = os.path.join(ROOT, "build", name)
    if not os.path.exists(p):
        lib = os.path.join(ROOT, "nbody_b200", "libnbody_b200.so")
        if not os.path.exists(lib):
            pytest.fail(f"{lib} missing: run __graft_entry__.build()")
        os.makedirs(os.path.join(ROOT, "build"), exist_ok=True)
        subprocess.run(["g++", "-std=c++20", "-O2", "-I" + os.path.join(ROOT, "include"), os.path.join(ROOT, _SOURCES[name]),
                        "-o", p, "-L" + os.path.join(ROOT, "nbody_b200"), "-lnbody_b200",
                        "-Wl,-rpath,$ORIGIN/../nbody_b200"], check=True)
    return p


def test_cpp_host_api_cases():
    r = subprocess.run([_exe("test_host_api")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "all checks passed" in r.stdout


def test_cpp_benchmark_main_record_shape():
    r = subprocess.run([_exe("benchmark_main"), "--particles", "8192", "--iterations", "3"], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0, r.stderr
    rec = json.loads(r.stdout)["benchmarks"]
    assert [b["benchmark_name"] for b in rec] == ["force.direct_n2", "force.barnes_hut", "force.spatial_hash",
                                                  "integration.direct_n2"]
    for b in rec:
        assert b["particle_count"] == 8192 and b["iterations"] == 3 and b["metrics"]["wall_time_ms"] > 0
