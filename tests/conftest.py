import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200, sm_100a)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(autouse=True)
def _ffb_env_overrides_reloaded():
    """The library caches its FFB_* override variables; tests that set one call lib.reload_env() right after, and this
    fixture (set up before monkeypatch, so torn down after it has restored the environment) reloads them again so the
    next test starts from the real environment."""
    yield
    try:
        from flexflow_b200 import lib
        if lib._lib is not None:
            lib.reload_env()
    except Exception:  # noqa: BLE001
        pass
