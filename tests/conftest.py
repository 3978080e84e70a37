import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu on the GPU box")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as O
    O.build()
    return O


@pytest.fixture(scope="session")
def fluid():
    """the product package; the CUDA library must already be built (no JIT, no fallback)."""
    import gpu_fluid_simulation_b200 as g
    from gpu_fluid_simulation_b200 import _build
    if _build.needs_build():
        _build.build()
    g.load()
    return g
