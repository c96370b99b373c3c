import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def ctx():
    """One rpvg_b200 context on cuda:0 for the whole GPU session (fails loudly if the .so or the GPU is missing)."""
    from rpvg_b200 import capi
    c = capi.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def port():
    from oracle import loader
    return loader.port()


@pytest.fixture(scope="session")
def reference():
    from oracle import loader
    if not loader.reference_available():
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    return loader.reference()
