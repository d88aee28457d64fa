import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session", autouse=True)
def _built_library():
    """Every test session works against the in-tree librnnwf.so; build it if it is missing/stale."""
    import __graft_entry__ as g
    g.build()


@pytest.fixture(scope="session")
def cuda():
    import torch
    if not torch.cuda.is_available():
        pytest.fail("test is marked gpu but no CUDA device is visible")
    torch.cuda.set_device(0)
    return torch.device("cuda", 0)


@pytest.fixture(autouse=True)
def _poison_device_memory(request):
    """GPU tests: fill the caching allocator's free memory with NaNs first, so that `torch.empty` workspaces / saved-state
    buffers hand the kernels garbage instead of the zeros a fresh process would see (an uninitialised read that only
    works on zeroed memory is a bug; this is how the padding-unit bug of the prefix-reuse path was found)."""
    if request.node.get_closest_marker("gpu") is None:
        yield
        return
    import torch
    if torch.cuda.is_available():
        junk = torch.full((3 << 27,), float("nan"), dtype=torch.float32, device="cuda")   # 1.5 GB
        del junk
    yield
