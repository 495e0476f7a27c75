import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(autouse=True)
def _fresh_oracle_state():
    """go_decompose keeps `hwidth` between calls (src/parallel_mod.f90:103-105); start clean."""
    from oracle import oracle
    oracle.reset_saved_state()
    yield
