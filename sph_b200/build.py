"""Builds the sm_100a C-ABI library (sph_b200/lib/libsphb200.so) with nvcc, in-tree.

The library has no torch / Python dependency: it is what the reference's C++
shim links.  nvcc cross-compiles without a GPU.
"""
from __future__ import annotations

import hashlib
import os
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
CSRC = ROOT / "csrc"
LIBDIR = ROOT / "lib"
LIB = LIBDIR / "libsphb200.so"

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O3",
    "-ccbin", "/usr/bin/g++",
    "--expt-relaxed-constexpr",
    "-cudart", "shared",
    "-Xlinker", "-rpath,/usr/local/cuda/lib64",
]


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def _digest() -> str:
    h = hashlib.sha256()
    for p in sorted(list(CSRC.glob("*.cu")) + list(CSRC.glob("*.cuh")) + [ROOT.parent / "include" / "sphb.h", Path(__file__)]):
        h.update(p.name.encode())
        h.update(p.read_bytes())
    return h.hexdigest()


def build(force: bool = False, verbose: bool = False, profile: bool = False, tag: str = "", defines: tuple = ()) -> Path:
    """profile=True builds lib/libsphb200_prof.so with the sweep's cycle counters compiled in
    (-DSPHB_KNN_PROFILE; select it with SPHB_LIB=<path>, enable with SPHB_KNN_PROF=1)."""
    LIBDIR.mkdir(exist_ok=True)
    tag = tag or ("_prof" if profile else "")
    defines = list(defines) + (["-DSPHB_KNN_PROFILE", "-DSPHB_KNN_TUNING"] if profile else [])
    lib = LIBDIR / f"libsphb200{tag}.so"
    stamp = LIBDIR / f"build{tag}.stamp"
    dig = _digest()
    if not force and lib.exists() and stamp.exists() and stamp.read_text() == dig:
        return lib
    objs = []
    procs = []
    for src in sources():
        obj = LIBDIR / (src.stem + tag + ".o")
        cmd = [NVCC, *FLAGS, *defines, "-c", str(src), "-o", str(obj)]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
        objs.append(obj)
    failed = False
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode != 0:
            sys.stderr.write(f"--- {src.name}\n{out}\n")
        failed |= p.returncode != 0
    if failed:
        raise RuntimeError("nvcc failed")
    cmd = [NVCC, *FLAGS, "-shared", "-o", str(lib), *map(str, objs), "-ldl"]
    subprocess.run(cmd, check=True)
    stamp.write_text(dig)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, profile="--profile" in sys.argv))
