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


@pytest.fixture(scope="session")
def b200_harness():
    """tests/harness/estimator_harness.cpp compiled against this repository's drop-in estimator classes."""
    from oracle import loader
    path = os.path.join(ROOT, "tests", "_build", "libharness_b200.so")
    if not os.path.exists(path):
        pytest.fail(f"{path} is missing: run __graft_entry__.build()")
    lib = loader.load_harness(path)
    assert lib.harness_impl_name() == b"rpvg_b200"
    return lib
