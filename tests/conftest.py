import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def oracle_layout(L):
    import qc_oracle as O
    return O.Layout(L.d, L.c, L.m, L.N, L.S, L.z_dim, L.x_off, L.u_off, L.dt_off, L.jac_mode)


def rel_err(a, b):
    """Block-normwise relative error (|a-b| <= tol * max|ref|)."""
    import numpy as np
    a, b = np.asarray(a), np.asarray(b)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


@pytest.fixture(scope="session")
def built_libs():
    """Make sure the product library and the oracle port are built (no compute)."""
    import subprocess
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True)
    lib = os.path.join(ROOT, "quantumcollocation.jl_b200", "csrc", "libqcb200.so")
    if not os.path.exists(lib):
        subprocess.run(["make", "-s", "-C", os.path.dirname(lib)], check=True)
    return lib
