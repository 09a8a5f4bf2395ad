import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
SEED = 0x4C696665415049


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def golden(name: str):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def port():
    from oracle.loader import Port
    return Port()


@pytest.fixture(scope="session")
def ref():
    """The unmodified reference compiled here (oracle/_ref); tests that need it skip if absent."""
    from oracle.loader import Ref
    r = Ref.load()
    if r is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    return r


@pytest.fixture(scope="session")
def engine():
    """The CUDA engine through the C ABI.  No fallback: without a GPU this raises."""
    from lifeapi_b200 import Engine
    e = Engine(0)
    yield e
    e.close()
