import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "orb-slam2_b200"), os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


CONFIGS = {
    # name: (width, height, nfeatures, scaleFactor, nlevels, iniThFAST, minThFAST) -- BASELINE.json configs
    "tum": (640, 480, 1000, 1.2, 8, 20, 7),
    "kitti": (1241, 376, 2000, 1.2, 8, 20, 7),
    "euroc": (752, 480, 1200, 1.2, 8, 20, 7),
}


@pytest.fixture(scope="session")
def configs():
    return CONFIGS
