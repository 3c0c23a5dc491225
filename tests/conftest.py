import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def lib_path():
    """librwkv.so, built in-tree (no-op when the stamp matches)."""
    from rwkv_cppy_b200.build import build_library
    return build_library(verbose=False)


@pytest.fixture(scope="session")
def raw_lib(lib_path):
    """The product library through the plain C-ABI binding (shared with the reference loader)."""
    from oracle.refload import RwkvLib
    L = RwkvLib(lib_path)
    return L


@pytest.fixture(scope="session")
def shared_library(lib_path):
    from rwkv_cppy_b200 import load_rwkv_shared_library
    return load_rwkv_shared_library()
