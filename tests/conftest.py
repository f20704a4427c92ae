import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _ensure_lib():
    so = os.path.join(ROOT, "bacchus-gpu-controller_b200", "libbgc.so")
    if not os.path.exists(so):
        import __graft_entry__ as g
        g.build()


@pytest.fixture(scope="session")
def bgc():
    """The product package (hyphenated directory name → importlib)."""
    _ensure_lib()
    return importlib.import_module("bacchus-gpu-controller_b200")


@pytest.fixture(scope="session")
def oracle():
    import oracle as o
    o.build()
    return o


@pytest.fixture()
def engine(bgc):
    import torch
    if not torch.cuda.is_available():
        pytest.fail("gpu-marked test running without a CUDA device: the product has no CPU fallback")
    e = bgc.Engine(0)
    yield e
    e.close()
