"""The INTEGRATION.md shims compiled as a C++ translation unit against include/sphb.h and linked with libsphb200.so only
(tests/shim/sph_shim.cpp): the boundary exercised from C++ through the reference's container shapes (utils::Graph arrays, a
MapMemEff stand-in), not only from Python."""
import subprocess
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent
SRC = ROOT / "tests" / "shim" / "sph_shim.cpp"
EXE = ROOT / "tests" / "shim" / "_build" / "sph_shim"


def _build():
    from sph_b200 import capi

    lib = capi.lib_path()
    EXE.parent.mkdir(exist_ok=True)
    cmd = ["/usr/bin/g++", "-std=c++17", "-O2", "-Wall", "-Wextra", "-Werror", str(SRC), "-o", str(EXE), f"-L{lib.parent}", "-lsphb200",
           f"-Wl,-rpath,{lib.parent}", "-Wl,-rpath,/usr/local/cuda/lib64"]
    subprocess.run(cmd, check=True, capture_output=True, text=True)
    return EXE


def test_shim_compiles_and_links_against_the_c_abi():
    exe = _build()
    ldd = subprocess.run(["ldd", str(exe)], capture_output=True, text=True).stdout
    assert "libsphb200.so" in ldd and "oracle" not in ldd and "torch" not in ldd


@pytest.mark.gpu
def test_shim_runs_on_the_device():
    exe = _build()
    r = subprocess.run([str(exe)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "shim ok" in r.stdout
