import os
import sys

import numpy as np
import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "oracle"))      # the oracle is test infrastructure
GOLDEN = os.path.join(REPO, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_routes(smooth=False):
    d = np.load(os.path.join(REPO, "mpc_for_av_at_intersection_b200", "data", "intersection_routes.npz"))
    routes = [d[f"route_{i}"].copy() for i in range(len(d["length"]))]
    if smooth:
        import mpc_oracle as mo
        for r in routes:
            mo.smooth_yaw(r[:, 2])
    return routes


@pytest.fixture(scope="session")
def routes_raw():
    return load_routes(False)


@pytest.fixture(scope="session")
def routes_smooth():
    return load_routes(True)


def golden(name):
    return np.load(os.path.join(GOLDEN, name))
