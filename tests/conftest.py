import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import binding
    return binding.oracle()


@pytest.fixture(scope="session")
def reference():
    """The reference's own compiled hot path; only exists where oracle/_ref was built (this container)."""
    from oracle import binding
    r = binding.reference()
    if r is None:
        pytest.skip("oracle/_ref/libtb200_ref.so not built (no /root/reference here)")
    return r


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    return np.load(ROOT / "tests" / "golden" / "hotpath_golden.npz")


@pytest.fixture(scope="session")
def tb():
    """The product, through its C ABI. GPU tests only: raises when the library or a B200 is missing."""
    from llamafile_b200.build import build
    build()
    import llamafile_b200 as tb
    if not tb.link():
        raise RuntimeError("no sm_100 device: the CUDA path cannot run and there is no fallback")
    return tb
