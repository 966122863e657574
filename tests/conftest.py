import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def product_lib():
    """The real C-ABI library (built by nvcc). Loading it needs no GPU; calling into it does."""
    from metta_b200.csrc import build as b

    b.build()
    from metta_b200 import capi

    return capi.lib()


@pytest.fixture(scope="session")
def emu_lib(product_lib):
    from emu_build import load_emu

    return load_emu()
