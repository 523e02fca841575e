import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def port():
    from oracle import pyoracle
    return pyoracle.Oracle("port")


@pytest.fixture(scope="session")
def ref():
    """The reference's own sources compiled as CPU code (oracle/_ref); built here when
    /root/reference exists, otherwise the prebuilt .so that travelled with the snapshot."""
    from oracle import pyoracle
    try:
        return pyoracle.Oracle("reference")
    except Exception as e:  # pragma: no cover
        pytest.skip(f"reference oracle unavailable: {e}")


@pytest.fixture(scope="session")
def vr():
    import volumetric_renderer_b200 as m
    return m
