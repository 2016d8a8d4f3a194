"""pytest configuration: the `gpu` marker and import paths.

`-m "not gpu"` runs here (no GPU): oracle vs golden vectors, host logic, C-ABI
symbol checks.  `-m gpu` runs on a B200 and calls the sm_100a kernels through
the C ABI, comparing against the oracle / the reference extension in oracle/_ref.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests"), os.path.join(ROOT, "gaussian-splatting-reflection_b200", "python")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:  # noqa: BLE001
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device in this container")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def lib_path():
    p = os.path.join(ROOT, "gaussian-splatting-reflection_b200", "lib", "libb200gs.so")
    if not os.path.exists(p):
        sys.path.insert(0, os.path.join(ROOT, "gaussian-splatting-reflection_b200"))
        import build as _build
        _build.build()
    return p
