import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")
    config.addinivalue_line("markers", "needs_ref: needs oracle/_ref/libsvtref.so built from /root/reference")


@pytest.fixture(scope="session")
def orc():
    from oracle import pyoracle
    pyoracle.build(ref=True)
    return pyoracle
