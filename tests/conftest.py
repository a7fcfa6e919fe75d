import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import kvoracle
    kvoracle.lib()  # builds on demand
    return kvoracle


@pytest.fixture(scope="session")
def kvlib():
    """libkvdp.so, built in-tree if needed. GPU tests additionally require a visible device (fail, never skip)."""
    from dynamo_exp_b200 import _lib, build
    build.build()
    return _lib


@pytest.fixture(scope="session")
def gpu(kvlib):
    kvlib.require_gpu()
    return kvlib
