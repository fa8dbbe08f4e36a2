import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (test infrastructure): the checker, never the thing under test."""
    import oracle as O

    return O


@pytest.fixture(scope="session")
def tex():
    """The product: ctypes mirror of include/rosy_tex_cuda.h over librosy_tex_cuda.so."""
    from rosy_b200 import texture

    texture.lib()  # fails loudly if the CUDA extension has not been built
    return texture


@pytest.fixture(scope="session")
def synth():
    from rosy_b200 import synth as S

    return S


def max_abs_diff(a, b):
    return int(np.abs(a.astype(np.int16) - b.astype(np.int16)).max()) if a.size else 0
