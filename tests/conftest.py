import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built_lib():
    """libtafl_b200.so built in-tree (nvcc cross-compiles without a GPU)."""
    from alphazeroforhnefatafl_b200 import build
    return build.build()


@pytest.fixture(scope="session")
def orc():
    import oracle_ffi
    return oracle_ffi.oracle()


@pytest.fixture(scope="session")
def sim():
    import oracle_ffi
    return oracle_ffi.hostsim()
