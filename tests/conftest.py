import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _load_ref(which):
    from oracle import build_ref
    try:
        return build_ref.load(which)
    except Exception as e:  # pragma: no cover - diagnostic only
        print(f"[conftest] could not load reference oracle {which}: {e}")
        return None


@pytest.fixture(scope="session")
def ref_fused():
    m = _load_ref("fused")
    if m is None:
        pytest.skip("oracle/_ref/_ref_fused.so not built (python oracle/build_ref.py fused)")
    return m


@pytest.fixture(scope="session")
def ref_sm80():
    m = _load_ref("sm80")
    if m is None:
        pytest.skip("oracle/_ref/_ref_sm80.so not built")
    return m


@pytest.fixture(scope="session")
def ref_sm89():
    m = _load_ref("sm89")
    if m is None:
        pytest.skip("oracle/_ref/_ref_sm89.so not built")
    return m


def accuracy(a, b):
    """(cosine similarity, relative L1, rmse) of a against b."""
    import torch
    a = a.float().flatten()
    b = b.float().flatten()
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    l1 = ((a - b).abs().sum() / b.abs().sum()).item()
    rmse = torch.sqrt(torch.mean((a - b) ** 2)).item()
    return cos, l1, rmse
