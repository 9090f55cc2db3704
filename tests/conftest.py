import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle as o
    o.build()
    return o


@pytest.fixture(scope="session")
def golden():
    import json
    here = os.path.join(ROOT, "tests", "golden")
    out = {}
    for name in ("reference_vectors", "synthetic_windows", "genesis_full_scan"):
        p = os.path.join(here, name + ".json")
        if os.path.exists(p):
            out[name] = json.load(open(p))
    return out


@pytest.fixture(scope="session")
def product():
    """The product package with its C-ABI library built (CPU-side build only)."""
    from otedama_b200 import build
    build.build()
    import otedama_b200
    return otedama_b200
