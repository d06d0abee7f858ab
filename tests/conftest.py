import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# the trunk modules raise instead of warning when a CUDA activation misses a library path
# (convnext._off_path); CPU tensors — the oracle side of the tests — are unaffected
os.environ.setdefault("CNX_STRICT", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def cnx_lib():
    """Build (if needed) and load libcnx_b200.so; nvcc cross-compiles without a GPU."""
    from image_classification_b200 import _lib, build
    build.build_library()
    return _lib.load()
