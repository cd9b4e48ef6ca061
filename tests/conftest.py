import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

REFERENCE_SRC = "/root/reference/src"


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu under gpurun)")


def gpu_available() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def native_lib():
    """Builds (if needed) and loads libflatip_b200.so -- works without a GPU (nvcc cross-compiles)."""
    import __graft_entry__ as entry
    from rag_inference_pipeline_b200 import _native

    if not os.path.exists(_native.LIB_PATH):
        entry.build()
    return _native.load_library()
