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
def golden():
    import json

    import numpy as np

    g = np.load(os.path.join(ROOT, "tests", "golden", "golden.npz"))
    with open(os.path.join(ROOT, "tests", "golden", "manifest.json")) as f:
        m = json.load(f)

    class G:
        manifest = m

        def __getitem__(self, k):
            return g[k].tobytes()

        def keys(self):
            return list(g.keys())

    return G()


@pytest.fixture(scope="session")
def oracle():
    import mrhelpers

    return mrhelpers.Oracle()
