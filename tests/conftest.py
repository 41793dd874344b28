import json
import os
import sys

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
for p in (HERE, ROOT):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run by the driver with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(HERE, "golden", "golden.json")) as f:
        return json.load(f)


def unhex(lst, dtype=np.float64):
    return np.array([float.fromhex(v) for v in lst], dtype=np.float64).astype(dtype)


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(HERE, "golden")
