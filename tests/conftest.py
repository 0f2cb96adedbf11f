import os
import sys

import pytest

REPO = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if REPO not in sys.path:
    sys.path.insert(0, REPO)
sys.path.insert(0, os.path.join(REPO, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with `-m gpu`)")


@pytest.fixture(scope="session")
def rtlib():
    """The product library, built in-tree (nvcc cross-compiles without a GPU)."""
    from ray_tracing_in_one_weekend_b200 import build
    build.build_all()
    import ray_tracing_in_one_weekend_b200 as rt
    rt.load_library()
    return rt


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    oracle_lib.lib()
    return oracle_lib
