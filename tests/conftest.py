import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build_oracle()
    return pyoracle.Oracle()


@pytest.fixture(scope="session")
def reflib():
    """The unmodified reference (oracle/_ref).  Built on demand where /root/reference exists;
    tests that need it are skipped elsewhere (the GPU box), where the committed golden fixtures
    stand in for it."""
    from oracle import pyoracle
    if not os.path.exists(pyoracle.REF_SO):
        if not os.path.isdir("/root/reference/native/src/seal"):
            pytest.skip("reference harness not built and /root/reference absent")
        pyoracle.build_ref()
    return pyoracle.RefLib()


@pytest.fixture(scope="session")
def simlib():
    """TEST-ONLY: the kernel sources compiled by g++ against tests/cusim (no GPU)."""
    sys.path.insert(0, os.path.join(ROOT, "tests", "cusim"))
    import build_sim
    from gemini_seal_b200 import _lib
    import ctypes
    return _lib.bind(ctypes.CDLL(build_sim.build()))


@pytest.fixture(scope="session")
def cudalib():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from gemini_seal_b200 import _lib
    return _lib.load()
