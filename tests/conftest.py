import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return dict(np.load(os.path.join(GOLDEN, name + ".npz")))


@pytest.fixture(scope="session")
def golden():
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = load_golden(name)
        return cache[name]
    return get


def jac_relerr(a, ref):
    """north_star metric for Jacobian entries: |d| <= tol * max(|ref|, 1e-2 * rowmax)
    (the floor absorbs the reference's 1e-100 epsilon artefacts)."""
    rowmax = np.abs(ref).max(axis=1, keepdims=True)
    scale = np.maximum(np.abs(ref), 1e-2 * rowmax)
    scale = np.where(scale == 0, 1.0, scale)
    return float((np.abs(a - ref) / scale).max())


def param_relerr(P, ref):
    """converged parameters: relative error with the scale floored at 1e-3 (true zeros such as
    the tangential terms); Euler angles compared modulo 360 degrees.  The skew gamma (index 2)
    is measured at the scale of alpha: it multiplies yd in u = alpha xd + gamma yd + uc exactly
    as alpha multiplies xd, its true value is 0 in every configuration, and past the cost
    plateau of a noisy run it is the weakest-determined entry (the reference's own explicit
    inverse moves it at the 1e-8 level, tests/test_oracle.py)."""
    P = np.asarray(P).ravel()
    ref = np.asarray(ref).ravel()
    d = np.abs(P - ref)
    d = np.minimum(d, np.abs(d - 360.0))
    scale = np.maximum(np.abs(ref), 1e-3)
    scale[2] = max(scale[2], abs(ref[0]))
    return float((d / scale).max())
