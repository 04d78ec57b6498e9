import os
import sys

import pytest

REPO = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
if REPO not in sys.path:
    sys.path.insert(0, REPO)

REFERENCE = "/root/reference"
SHIM = os.path.join(REPO, "tests", "golden", "_make")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (build container only)")


def have_reference():
    return os.path.isdir(os.path.join(REFERENCE, "evostencils"))


@pytest.fixture(scope="session")
def reference_front_end():
    """The unmodified reference grammar/IR/emitter, importable through the deap shim."""
    if not have_reference():
        pytest.skip("reference checkout not present")
    for p in (SHIM, REFERENCE):
        if p not in sys.path:
            sys.path.insert(0, p)
    sys.path.insert(0, SHIM)
    import make_lowering_golden as m
    return m
