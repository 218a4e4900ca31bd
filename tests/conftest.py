import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Build the in-tree artefacts once if they are missing (no-op on the GPU box, where the
    prebuilt .so files travel with the snapshot)."""
    from pine_b200 import _abi
    port = os.path.join(ROOT, "oracle", "liboracle_port.so")
    tables = os.path.join(ROOT, "pine_b200", "data", "bluesobol_tables.bin")
    if not (os.path.exists(_abi.LIB_PATH) and os.path.exists(port) and os.path.exists(tables)):
        import __graft_entry__
        __graft_entry__.build()
    yield


def _has_ref():
    return os.path.exists(os.path.join(ROOT, "oracle", "_ref", "libpine_ref.so"))


needs_ref = pytest.mark.skipif(not _has_ref(), reason="oracle/_ref (compiled reference) not built")


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    cache = {}

    def load(name):
        if name not in cache:
            cache[name] = dict(np.load(os.path.join(ROOT, "tests", "golden", name + ".npz")))
        return cache[name]
    return load
