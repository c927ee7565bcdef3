"""TEST INFRASTRUCTURE ONLY — builds tests/emu/_build/libsphb200_emu.so: the repo's CSR / graph kernels (csrc/*.cu, the
product sources, unmodified) compiled as plain C++ against the CPU stand-in of the CUDA runtime in tests/emu/include.
The only source transformation is syntactic: `kernel<<<grid, block, smem, stream>>>(args);` becomes
`emu::launch(grid, block, smem, [&] { kernel(args); });` and `extern __shared__ T name[];` becomes a pointer to the block's
dynamic shared buffer.  The tensor-core kNN sources (tcgen05 / TMA inline PTX) are not part of this build."""
from __future__ import annotations

import re
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
ROOT = HERE.parent.parent
CSRC = ROOT / "sph_b200" / "csrc"
OUT = HERE / "_build"
LIB = OUT / "libsphb200_emu.so"
SOURCES = ["graph_merge.cu", "graph_ops.cu", "level_ops.cu", "sims.cu", "walks.cu", "prow.cu", "geodesic.cu", "level_api.cu", "cache_io.cu"]


def _split_top(s: str) -> list[str]:
    parts, depth, cur = [], 0, []
    for ch in s:
        if ch in "([{":
            depth += 1
        elif ch in ")]}":
            depth -= 1
        if ch == "," and depth == 0:
            parts.append("".join(cur).strip())
            cur = []
        else:
            cur.append(ch)
    parts.append("".join(cur).strip())
    return parts


def translate(text: str) -> str:
    text = re.sub(r"extern\s+__shared__\s+(\w+)\s+(\w+)\s*\[\s*\]\s*;", r"\1* \2 = reinterpret_cast<\1*>(emu::dyn_smem());", text)
    out, pos = [], 0
    while True:
        i = text.find("<<<", pos)
        if i < 0:
            out.append(text[pos:])
            break
        m = re.search(r"([A-Za-z_]\w*(?:<[^<>()]*>)?)\s*$", text[pos:i])
        assert m, text[max(0, i - 80):i + 20]
        name_start = pos + m.start(1)
        j = text.index(">>>", i)
        cfg = _split_top(text[i + 3:j])
        assert len(cfg) >= 2, cfg
        k = j + 3
        while text[k].isspace():
            k += 1
        assert text[k] == "(", text[k:k + 40]
        depth, e = 0, k
        while True:
            if text[e] == "(":
                depth += 1
            elif text[e] == ")":
                depth -= 1
                if depth == 0:
                    break
            e += 1
        args = text[k + 1:e]
        smem = cfg[2] if len(cfg) > 2 else "0"
        out.append(text[pos:name_start])
        out.append(f"emu::launch({cfg[0]}, {cfg[1]}, {smem}, [&] {{ {m.group(1)}({args}); }})")
        pos = e + 1
    return "".join(out)


def build(verbose: bool = False, asan: bool = False, defines: tuple = (), tag: str = "") -> Path:
    """asan=True builds libsphb200_emu_asan.so with AddressSanitizer: every "device" buffer is a heap block, so an
    out-of-bounds global-memory access of a kernel is reported (run with LD_PRELOAD=$(gcc -print-file-name=libasan.so)
    ASAN_OPTIONS=detect_leaks=0:detect_stack_use_after_return=0)."""
    OUT.mkdir(exist_ok=True)
    lib = OUT / "libsphb200_emu_asan.so" if asan else (OUT / f"libsphb200_emu{tag}.so" if tag else LIB)
    srcs = []
    for name in SOURCES:
        dst = OUT / (Path(name).stem + "_emu.cpp")
        dst.write_text(f'#line 1 "{CSRC / name}"\n' + translate((CSRC / name).read_text()))
        srcs.append(dst)
    cmd = ["/usr/bin/g++", "-O1", "-g", "-std=c++17", "-fPIC", "-shared", "-ffp-contract=off", "-fvisibility=hidden", "-D__CUDACC__",
           *(["-fsanitize=address", "-fno-omit-frame-pointer"] if asan else []), *defines,
           "-w", f"-I{HERE / 'include'}", f"-I{CSRC}", "-o", str(lib), *map(str, srcs), str(HERE / "emu_api.cpp"), str(HERE / "emu_stubs.cpp")]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0 or verbose:
        sys.stderr.write(r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("emulation build failed")
    return lib


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, asan="--asan" in sys.argv))
