import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

RES = 0.05
FOOTPRINTS = {
    "rect_hole": ([0.64, 0.46], True),
    "rect_solid": ([0.64, 0.46], False),
    "box_hole": ([-0.2, -0.25, 0.6, 0.25], True),
    "circle_hole": ([0.35], True),
}
OFFSETS = {"centre": (0.5, 0.5, 0.0), "corner": (0.0, 0.0, 0.0), "frac": (0.37, 0.81, 0.29)}


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    return np.load(os.path.join(ROOT, "tests", "golden", "hotpath_v1.npz"))


@pytest.fixture(scope="session")
def port():
    import oracle
    return oracle.load_port()


@pytest.fixture(scope="session")
def ref():
    """The reference's own code; None on machines where oracle/_ref was never built."""
    import oracle
    return oracle.load_ref()


@pytest.fixture()
def handle():
    from path_planner_b200 import capi
    h = capi.Handle(0)
    yield h
    h.close()
