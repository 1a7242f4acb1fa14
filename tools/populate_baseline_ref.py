#!/usr/bin/env python
"""Populate baseline/_ref/ (git-ignored, travels to the GPU box with gpurun) with the few UNMODIFIED reference Python
files the drop-in test needs: the two `gaussian_renderer/__init__.py` (the caller of the rasterizer this repo replaces:
gs3dgs/gaussian_renderer/__init__.py:24-133, gs2dgs/gaussian_renderer/__init__.py:24-160) and the two helpers they
import that are plain torch (utils/sh_utils.py, utils/point_utils.py). Nothing is copied into the tracked tree and
nothing under scorp_b200/ reads baseline/_ref; tests/test_gpu_reference_render.py executes these files on top of
scorp_b200.install() and on top of the reference CUDA, and compares.

    python tools/populate_baseline_ref.py [--ref /root/reference]

__graft_entry__.build() runs it when the reference tree is present (this container); the GPU box only uses the copies.
"""
from __future__ import annotations

import argparse
import shutil
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
FILES = [f"{pkg}/{rel}" for pkg in ("gs3dgs", "gs2dgs")
         for rel in ("gaussian_renderer/__init__.py", "utils/sh_utils.py", "utils/point_utils.py")]


def populate(ref: Path = Path("/root/reference"), dst: Path = ROOT / "baseline" / "_ref") -> int:
    n = 0
    for rel in FILES:
        src = ref / rel
        if not src.is_file():
            continue
        out = dst / rel
        out.parent.mkdir(parents=True, exist_ok=True)
        shutil.copyfile(src, out)
        n += 1
    return n


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--ref", default="/root/reference")
    a = ap.parse_args()
    print(f"copied {populate(Path(a.ref))} reference files into {ROOT / 'baseline' / '_ref'}")
