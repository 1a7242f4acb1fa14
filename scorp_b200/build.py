"""In-tree build of libgsr_b200.so (hand-written sm_100a CUDA behind a C ABI).

No torch headers are involved: the library is plain CUDA C++ compiled by nvcc, so a full
rebuild takes seconds. The shared object is written next to this file (git-ignored, shipped
to the GPU box by gpurun).
"""
from __future__ import annotations

import os
import subprocess
import sys
from pathlib import Path

HERE = Path(__file__).resolve().parent
CSRC = HERE / "csrc"
LIB = HERE / "libgsr_b200.so"
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

# Default -fmad (contraction on) and no fast-math: same floating-point model as the reference
# build. -lineinfo keeps ncu's source page usable.
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo", "-shared",
    "-Xcompiler", "-fPIC,-O2,-Wall",
    "-Xptxas", "-v",
    "-ldl",          # nvtx3 (header-only) loads its injection library on demand
]


def sources() -> list[Path]:
    return sorted(CSRC.glob("*.cu"))


def needs_build() -> bool:
    if not LIB.exists():
        return True
    t = LIB.stat().st_mtime
    deps = list(CSRC.glob("*")) + [HERE.parent / "include" / "gsr_b200.h", Path(__file__)]
    return any(p.stat().st_mtime > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> Path:
    if not force and not needs_build():
        return LIB
    cmd = [NVCC, *FLAGS, *map(str, sources()), "-o", str(LIB)]
    proc = subprocess.run(cmd, capture_output=True, text=True)
    log = proc.stdout + proc.stderr
    (HERE / "build.log").write_text(" ".join(cmd) + "\n" + log)
    if proc.returncode != 0:
        sys.stderr.write(log)
        raise RuntimeError("nvcc failed building libgsr_b200.so (see scorp_b200/build.log)")
    if verbose:
        print(log)
    return LIB


if __name__ == "__main__":
    build(force=True, verbose=True)
    print("built", LIB)
