import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "sgl-kernel-xpu_b200")):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# never let pytest repr a (possibly faulted) device tensor — reference tests/conftest.py:4-10
torch.set_printoptions(threshold=0)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def quant_golden():
    return np.load(os.path.join(GOLDEN, "quant_golden.npz"))


@pytest.fixture(scope="session")
def gemm_golden():
    return np.load(os.path.join(GOLDEN, "gemm_golden.npz"))


@pytest.fixture(scope="session")
def grouped_golden():
    return np.load(os.path.join(GOLDEN, "grouped_golden.npz"))


@pytest.fixture(scope="session")
def token_golden():
    return np.load(os.path.join(GOLDEN, "token_golden.npz"))


@pytest.fixture(autouse=True)
def _reset_torch_state():
    # reference tests/conftest.py:16-37: default dtype/device reset + device sync around each test
    torch.set_default_dtype(torch.float32)
    yield
    if torch.cuda.is_available():
        torch.cuda.synchronize()


def from_bits(a: np.ndarray, dtype: torch.dtype) -> torch.Tensor:
    """Rebuild a tensor from the raw bit patterns stored in the golden files."""
    t = torch.from_numpy(np.ascontiguousarray(a))
    if dtype in (torch.bfloat16, torch.float16):
        return t.view(torch.int16).view(dtype)
    if dtype == torch.float8_e4m3fn:
        return t.view(torch.float8_e4m3fn)
    return t
