"""pytest configuration: `gpu` marker, import helpers for the hyphenated package and the oracle."""
import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

PACKAGE = "distributed-radxi-hash-join-on-gpus_b200"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def rhj():
    """The product package (ctypes binding of librhj_b200.so); builds the library if stale."""
    mod = importlib.import_module(PACKAGE)
    mod.build()
    mod.lib()
    return mod


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure)."""
    from oracle import oracle as O
    O.lib()
    return O
