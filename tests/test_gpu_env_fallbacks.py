"""The round-2a kernels stay in the library behind environment switches (the four- / eight-warp
CTA blends, the per-tile sort without its bucket pass, binning aggregated per CTA for a single view). The switches are read
once per process, so each combination runs a few of the parity cases in a process of its own: every kernel the library
can be told to launch is held to the same bit-exact / rtol 1e-4 bar as the default ones."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent

COMBOS = [
    ("cta_blends", {"GSR_FWD_WARP": "0", "GSR_BWD_WARP": "0"},
     [["tests/test_gpu_dgr_parity.py", "-k", "tiny_ragged or C2_100k or dense_mid"], ["tests/test_gpu_dsr_parity.py"]]),
    ("surfel_forward_one_warp", {"GSR_SURFEL_FWD_WARP": "1"}, [["tests/test_gpu_dsr_parity.py"]]),
    ("sort_without_bucket_pass_aggregated_binning", {"GSR_BUCKET_SORT": "0", "GSR_AGGREGATE": "1"},
     [["tests/test_gpu_dgr_parity.py", "-k", "dense_mid or dense_long or C2_100k or equal_depths or clustered"]]),
]


@pytest.mark.parametrize("name,env,runs", COMBOS, ids=[c[0] for c in COMBOS])
def test_parity_with_the_fallback_kernels(name, env, runs):
    e = dict(os.environ); e.update(env)
    for args in runs:
        p = subprocess.run([sys.executable, "-m", "pytest", "-m", "gpu", "-x", "-q", "-p", "no:cacheprovider", *args],
                           cwd=ROOT, env=e, capture_output=True, text=True, timeout=900)
        tail = "\n".join((p.stdout + p.stderr).splitlines()[-15:])
        assert p.returncode == 0, tail
        assert " passed" in p.stdout and " failed" not in p.stdout, tail
