import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import oracle

    oracle.build()
    return oracle.library(False)


@pytest.fixture(scope="session")
def product_lib():
    """The CUDA library through its C ABI; fails loudly if it is not built or no GPU is present."""
    from cr_mech_coli_b200 import _capi

    lib = _capi.product()
    n = lib.fn("device_count")()
    assert n > 0, "gpu tests need a CUDA device: " + lib.last_error()
    return lib
