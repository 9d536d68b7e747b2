import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference_dyn():
    """the unmodified reference engines (oracle/_ref), or skip when they were not built / host lacks AVX-512"""
    from oracle import pyoracle
    try:
        return pyoracle.Reference(None)
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference build unavailable: {e}")


@pytest.fixture(scope="session")
def reference_128():
    from oracle import pyoracle
    try:
        return pyoracle.Reference(128)
    except Exception as e:  # noqa: BLE001
        pytest.skip(f"reference build unavailable: {e}")
