import os
import sys

import numpy as np
import pytest

ROOT = os.path.normpath(os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)   # test-side helpers (bcnenc_host.py)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def golden_decode():
    return np.load(os.path.join(GOLDEN, "bc_decode.npz"))


@pytest.fixture(scope="session")
def golden_colour():
    return np.load(os.path.join(GOLDEN, "colour.npz"))


@pytest.fixture(scope="session")
def golden_pipeline():
    return np.load(os.path.join(GOLDEN, "pipeline.npz"))


@pytest.fixture(scope="session")
def golden_pipeline2():
    return np.load(os.path.join(GOLDEN, "pipeline2.npz"))


@pytest.fixture(scope="session")
def oracle():
    """CPU restatement (oracle/vq_oracle.cpp), built on demand with g++."""
    from oracle import orcbind
    orcbind.lib()
    return orcbind


@pytest.fixture(scope="session")
def reflib():
    """The unmodified reference (oracle/_ref), when it was built where /root/reference exists."""
    from oracle import refbind
    if not refbind.available():
        pytest.skip("oracle/_ref/libyupref.so not built (needs /root/reference)")
    refbind.lib()
    return refbind


def bits_equal(a, b):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b)
    return a.shape == b.shape and a.dtype == b.dtype and np.array_equal(a.view(np.uint8), b.view(np.uint8))
