import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """The libraries are built in-tree by __graft_entry__.build(); build them if a fresh checkout
    has not yet (CPU-only cross-compile, about a minute)."""
    need = [os.path.join(ROOT, "ray-tracer_b200", "librtb.so"), os.path.join(ROOT, "ray-tracer_b200", "librtb_scenes.so"),
            os.path.join(ROOT, "oracle", "libpt_oracle.so")]
    if not all(os.path.exists(p) for p in need):
        import __graft_entry__
        __graft_entry__.build()
