"""Import-name shim: `import diff_surfel_rasterization` -> scorp_b200's sm_100a implementation."""
from scorp_b200.diff_surfel_rasterization import *  # noqa: F401,F403
from scorp_b200.diff_surfel_rasterization import _C, _RasterizeGaussians  # noqa: F401
