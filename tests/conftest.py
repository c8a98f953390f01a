import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    import numpy as np

    cache = {}

    def load(name):
        if name not in cache:
            path = os.path.join(ROOT, "tests", "golden", name + ".npz")
            if not os.path.exists(path):
                pytest.skip("fixture %s.npz has not been generated (tests/golden/make_golden.py %s)" % (name, name))
            cache[name] = dict(np.load(path))
        return cache[name]

    return load
