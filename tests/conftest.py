import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
REFERENCE = "/root/reference"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden_dir():
    return GOLDEN


@pytest.fixture(scope="session")
def reference_aug():
    """The reference's own src.augmentations, imported read-only (build container only)."""
    if not os.path.isdir(os.path.join(REFERENCE, "src", "augmentations")):
        pytest.skip("reference tree not mounted (GPU box)")
    import importlib
    import io
    import contextlib
    sys.path.insert(0, REFERENCE)
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            return importlib.import_module("src.augmentations")
    finally:
        sys.path.remove(REFERENCE)
