import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = ["cfgS_cond", "uncond_small", "cfgA_mini", "cfgB_mini"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    """-> (sd: state_dict as numpy, data: the other arrays)."""
    raw = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    sd = {k[4:]: raw[k] for k in raw.files if k.startswith("sd::")}
    data = {k: raw[k] for k in raw.files if not k.startswith("sd::")}
    return sd, data


def rel_l2(a, b):
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(b), 1e-30))


def max_rel(a, b):
    """max |a-b| / max|b|  (the 'relative' of the north star: error relative to the tensor's scale)."""
    a = np.asarray(a, dtype=np.float64).ravel()
    b = np.asarray(b, dtype=np.float64).ravel()
    return float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30))


@pytest.fixture(params=GOLDEN_CASES)
def golden(request):
    sd, data = load_golden(request.param)
    return request.param, sd, data
