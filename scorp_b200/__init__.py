"""scorp_b200 -- B200-native (sm_100a) differentiable Gaussian-splatting rasterizer.

Drop-in replacements for the three CUDA extensions behind SCORP's gaussian_renderer.render():
  scorp_b200.diff_gaussian_rasterization   (3DGS)
  scorp_b200.diff_surfel_rasterization     (2DGS surfels)
  scorp_b200.simple_knn                    (distCUDA2)
`scorp_b200.install()` puts them on sys.path under the reference's own import names.
All compute is hand-written CUDA in libgsr_b200.so behind the C ABI of include/gsr_b200.h.
"""
from __future__ import annotations

import sys
from pathlib import Path

__version__ = "0.1.0"

_DROPIN = Path(__file__).resolve().parent / "dropin"


def install() -> None:
    """Make `import diff_gaussian_rasterization` (etc.) resolve to this package's drop-ins."""
    p = str(_DROPIN)
    if p not in sys.path:
        sys.path.insert(0, p)
