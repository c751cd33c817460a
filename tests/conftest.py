import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

DATA = os.path.join(ROOT, "tests", "data")
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def golden(name):
    with open(os.path.join(GOLDEN, name)) as f:
        return json.load(f)


def data_path(name):
    return os.path.join(DATA, name)


@pytest.fixture(scope="session")
def lib_built():
    """Builds libpdgpu.so if it is missing (nvcc cross-compiles without a GPU)."""
    from prairiedog_b200 import build
    build.build_library()
    return build.LIB
