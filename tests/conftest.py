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
