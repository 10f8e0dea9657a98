"""GPU: the drop-in claim, end to end.  oracle/_ref/ref_bench_on_b200 is the REFERENCE's own
benchmarks/benchmark_main.cpp + src/core/particle_system.cpp (unmodified sources) linked against
integration/b200_backend.cpp and libnbody_b200.so instead of the reference's four hot-path .cu files
(integration/build_dropin.sh).  It must run all five of the reference's benchmarks."""
import json
import os
import subprocess

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "oracle", "_ref", "ref_bench_on_b200")


def test_reference_benchmark_program_runs_on_our_backend(tmp_path):
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/ref_bench_on_b200 not built (needs /root/reference at build time)")
    out = tmp_path / "bench.json"
    r = subprocess.run([EXE, "--benchmark", "all", "--particle-count", "20000", "--iterations", "3", "--output",
                        str(out)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    rec = json.loads(out.read_text())["benchmarks"]
    names = [b["benchmark_name"] for b in rec]
    assert names == ["serialization.round_trip", "force.direct_n2", "force.barnes_hut", "force.spatial_hash",
                     "integration.velocity_verlet"]
    for b in rec:
        assert b["particle_count"] == 20000 and b["metrics"]["wall_time_ms"] >= 0
