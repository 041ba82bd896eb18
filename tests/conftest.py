import importlib
import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for it in items:
        if "gpu" in it.keywords:
            it.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    from _oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The compiled, unmodified reference (oracle/_ref). Skips when it is neither built nor buildable."""
    from _oracle import Reference
    try:
        return Reference()
    except FileNotFoundError as e:
        pytest.skip(str(e))


@pytest.fixture(scope="session")
def nv():
    """The product package. Fails loudly (no fallback) when the CUDA library is not built."""
    lib = ROOT / "nano-vectordb_b200" / "lib" / "libnvdb_b200.so"
    if not lib.exists():
        import subprocess
        subprocess.run(["make", "-C", str(ROOT), "lib"], check=True, capture_output=True)
    return importlib.import_module("nano-vectordb_b200")
