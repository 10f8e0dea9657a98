import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    from oracle import bindings
    bindings.lib()
    return bindings


def rel_err(a, ref):
    """max_i |a_i - ref_i|_2 / |ref_i|_2 over bodies; a, ref = tuples of 3 arrays."""
    import numpy as np
    d = np.sqrt(sum((np.asarray(x, np.float64) - np.asarray(y, np.float64)) ** 2 for x, y in zip(a, ref)))
    n = np.sqrt(sum(np.asarray(y, np.float64) ** 2 for y in ref))
    return d / np.maximum(n, 1e-300)
