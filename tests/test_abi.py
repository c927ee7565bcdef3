"""CPU tests (no GPU): the C-ABI library loads, exports every symbol include/sphb.h declares,
the ctypes table covers the header, and compute calls fail loudly without a device."""
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    text = (ROOT / "include" / "sphb.h").read_text()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"SPHB_API\s+[\w\s\*]+?\b(sphb_\w+)\s*\(", text)))


def test_header_symbols_exported(capi):
    names = _declared()
    assert len(names) >= 20
    out = subprocess.run(["nm", "-D", "--defined-only", str(capi.lib_path())], capture_output=True, text=True, check=True).stdout
    exported = set(re.findall(r"\bT\s+(sphb_\w+)", out))
    missing = [n for n in names if n not in exported]
    assert not missing, f"declared in sphb.h but not exported: {missing}"


def test_ctypes_table_matches_header(capi):
    assert sorted(capi.SIGNATURES) == _declared()
    lib = capi.load()
    for n in capi.SIGNATURES:
        assert getattr(lib, n) is not None


def test_no_torch_types_in_header():
    text = (ROOT / "include" / "sphb.h").read_text()
    assert "torch" not in text and "at::" not in text and "std::" not in text


def test_library_has_no_oracle_or_torch_dependency(capi):
    out = subprocess.run(["ldd", str(capi.lib_path())], capture_output=True, text=True).stdout
    assert "sph_oracle" not in out and "libtorch" not in out and "libpython" not in out


def test_sm100a_sass_present(capi):
    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", str(capi.lib_path())], capture_output=True, text=True).stdout
    assert "sm_100a" in out


def test_compute_fails_loudly_without_device(capi):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    x = np.zeros((64, 8), np.float32)
    with pytest.raises(capi.SphbError) as e:
        capi.knn_exact(x, 4)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)
    with pytest.raises(capi.SphbError):
        capi.init(0)


def test_product_never_imports_oracle():
    # the product path may not import, link or execute anything under oracle/
    for p in list((ROOT / "sph_b200").rglob("*.py")) + list((ROOT / "sph_b200" / "csrc").glob("*")) + [ROOT / "include" / "sphb.h"]:
        if p.is_file():
            t = p.read_text(errors="ignore")
            assert "oracle_lib" not in t and "libsph_oracle" not in t and "orc_" not in t, p
