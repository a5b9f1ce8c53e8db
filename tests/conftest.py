"""pytest configuration: markers, import paths, shared fixtures."""
import json
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _pin_nccl_to_torchs_copy():
    """libgpuprobe dlopen()s NCCL by soname. If it loads the system libnccl.so.2 first, a later `import torch`
    in the same process binds libtorch_cuda.so to that (older) copy and fails on missing symbols. Point the
    library at the copy torch ships, so that both use one NCCL whatever the import order."""
    if os.environ.get("GPUPROBE_NCCL_LIB"):
        return
    import importlib.util

    spec = importlib.util.find_spec("nvidia.nccl")
    for base in (spec.submodule_search_locations if spec and spec.submodule_search_locations else []):
        cand = os.path.join(base, "lib", "libnccl.so.2")
        if os.path.exists(cand):
            os.environ["GPUPROBE_NCCL_LIB"] = cand
            return


_pin_nccl_to_torchs_copy()


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
