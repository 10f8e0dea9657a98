"""GPU: the C++20 host layer (include/nbody/*.hpp) -- the reference's test cases re-expressed in C++
against the drop-in headers, and the benchmark program in the reference's JSON record shape."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _exe(name):
    p = os.path.join(ROOT, "build", name)
    if not os.path.exists(p):
        pytest.fail(f"{p} missing: run __graft_entry__.build()")
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
