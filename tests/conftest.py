import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    # the oracle (plain C) is test infrastructure: build it on demand, here and on the GPU box
    if not os.path.exists(os.path.join(ROOT, "oracle", "libedison_oracle.so")):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle"), "all"])


@pytest.fixture(scope="session")
def oracle():
    import common
    return common.Oracle()


@pytest.fixture(scope="session")
def edison():
    import edison_b200
    edison_b200.init()
    return edison_b200
