import os
import subprocess
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, str(ROOT / "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def _has_gpu() -> bool:
    if os.environ.get("SPHB_EMULATE") == "1":
        # tests/emu: the -m gpu tests of the CSR / graph passes run against the CPU emulation build of the same kernel
        # sources (SPHB_LIB must point at tests/emu/_build/libsphb200_emu.so); a logic check, never a measurement
        return True
    try:
        import torch

        return torch.cuda.is_available()
    except Exception:
        return False


def pytest_collection_modifyitems(config, items):
    if _has_gpu():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib

    return oracle_lib.load()


@pytest.fixture(scope="session")
def capi():
    from sph_b200 import capi as c

    c.load()
    return c
