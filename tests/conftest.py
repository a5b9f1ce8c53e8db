"""pytest configuration: markers, import paths, shared fixtures."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu under gpurun)")
    config.addinivalue_line("markers", "multigpu: needs at least 2 GPUs (subset of gpu)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "golden_v1.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def gp():
    """The ctypes mirror of include/gpuprobe.h (loads libgpuprobe.so; fails loudly if absent)."""
    import __graft_entry__ as ge

    return ge.load_package().gpuprobe
