import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def emul():
    """Host layer on the CPU-emulated device ABI (tests only)."""
    from tests.support import emulated_host_library
    return emulated_host_library()


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(ROOT, "tests", "golden", "amplitudes.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def gpu_lib():
    import qflex_b200
    return qflex_b200.default_library()
