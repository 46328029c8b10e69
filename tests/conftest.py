import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_report_header(config):
    try:
        import cv2
        return "oracle: cv2 %s (expected 4.13.0)" % cv2.__version__
    except Exception as e:  # pragma: no cover
        return "oracle: cv2 unavailable (%s)" % e


@pytest.fixture(scope="session")
def cv2ref():
    from oracle import cv2ref as m
    if not m.version_ok():
        pytest.skip("cv2 %s is not the pinned oracle version" % m.cv2.__version__)
    return m


@pytest.fixture(scope="session")
def port():
    from oracle import port as m
    m.build()
    return m
