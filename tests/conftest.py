import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

MODELS_DIR = os.path.join(ROOT, "tksm_b200", "models")
os.environ.setdefault("TKSM_MODELS", MODELS_DIR)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: long-running CPU test")


@pytest.fixture(scope="session")
def built():
    from tksm_b200 import _lib
    _lib.build()
    from oracle import seq_oracle
    seq_oracle.build(("libtwin.so", "libnw.so"))
    return True
