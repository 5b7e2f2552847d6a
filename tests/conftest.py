import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def scenes():
    import numpy as np
    out = {}
    for i, name in enumerate(("test1", "test13", "test47", "test56", "test60")):
        d = np.load(os.path.join(ROOT, "tests", "golden", "scene_%s.npz" % name))
        out[name] = dict(id=i, xyz=d["xyz"], rgba=d["rgba"], label=d["label"])
    return out
