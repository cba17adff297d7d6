import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


def cases_of(z, prefix_filter=None):
    """Group 'case/key' npz entries into {case: {key: array}}."""
    out = {}
    for k in z.files:
        if "/" not in k:
            continue
        case, key = k.rsplit("/", 1)
        if prefix_filter is None or prefix_filter(case):
            out.setdefault(case, {})[key] = z[k]
    return out


@pytest.fixture(scope="session")
def golden():
    return load_golden


# north_star's parity bars, used verbatim by the tests: relative error of smoothed means and
# variances <= 1e-5 in fp64 and <= 1e-3 in fp32.
RTOL_F64 = 1e-5
RTOL_F32 = 1e-3


def rel(a, b, time_axis=None):
    """Largest error of `a` against the expected `b`, each COMPONENT scaled by its own largest magnitude over time
    (so a small covariance entry is not graded against the largest number in the array).  time_axis defaults to 0
    for (T, ...) arrays; pass 1 for (B, T, ...).  A component that is (numerically) zero throughout -- the reference's
    all-zero S[T-1], a structurally zero covariance entry -- is scaled by 1e-6 of the array's largest magnitude."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape, (a.shape, b.shape)
    if b.size == 0:
        return 0.0
    if b.ndim == 0 or (b.ndim == 1 and time_axis is None):
        return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))
    ax = 0 if time_axis is None else time_axis
    scale = np.abs(b).max(axis=ax, keepdims=True)
    floor = max(float(np.abs(b).max()) * 1e-6, 1e-300)
    return float((np.abs(a - b) / np.maximum(scale, floor)).max())
