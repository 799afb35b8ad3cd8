import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """make sure the in-tree shared libraries exist (no-op when they are already built)"""
    import __graft_entry__ as g
    g.build(only_missing=True)


@pytest.fixture(scope="session")
def ctx():
    from ipk_b200 import capi as libipkgpu
    c = libipkgpu.Context(0)
    yield c
    c.close()
