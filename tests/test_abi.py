"""CPU-only: the C-ABI library loads and exports every symbol include/gsr_b200.h declares; the Python
boundary mirrors the reference's names and argument validation; the product never imports oracle/."""
import ctypes
import re
from pathlib import Path

import pytest
import torch

ROOT = Path(__file__).resolve().parent.parent


def _declared():
    txt = (ROOT / "include" / "gsr_b200.h").read_text()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(gsr_[a-z0-9_]+)\s*\(", txt)) - {"gsr_alloc_fn"})


def test_library_exports_every_declared_symbol():
    from scorp_b200 import _lib
    lib = _lib.lib()
    names = _declared()
    assert len(names) >= 18
    for n in names:
        assert hasattr(lib, n), n
        assert n in _lib.SIGNATURES, f"{n} declared in the header but not bound in scorp_b200/_lib.py"
    assert lib.gsr_version() >= 100


def test_layout_queries_are_pure_host_calls():
    from scorp_b200 import _lib
    lib = _lib.lib()
    off = (ctypes.c_size_t * 6)(); tot = ctypes.c_size_t(0)
    lib.gsr_geom_layout(1000, 64, off, ctypes.byref(tot))
    assert off[0] == 0 and all(o % 256 == 0 for o in off) and tot.value >= 1000 * 48 + 1000 * 24
    assert lib.gsr_backward_workspace_bytes(1000) >= 1000 * 40
    assert lib.gsr_knn_workspace_bytes(1000) > 0


def test_python_surface_matches_reference_names():
    import scorp_b200.diff_gaussian_rasterization as dgr
    import scorp_b200.diff_surfel_rasterization as dsr
    from scorp_b200.simple_knn._C import distCUDA2
    fields = ("image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
              "projmatrix", "sh_degree", "campos", "prefiltered", "debug")
    assert dgr.GaussianRasterizationSettings._fields == fields
    assert dsr.GaussianRasterizationSettings._fields == fields
    for mod in (dgr, dsr):
        r = mod.GaussianRasterizer(None)
        assert hasattr(r, "markVisible") and hasattr(r, "forward")
        with pytest.raises(Exception, match="SHs or precomputed colors"):
            r(torch.zeros(2, 3), torch.zeros(2, 3), torch.zeros(2, 1))
        with pytest.raises(Exception, match="scale/rotation pair"):
            r(torch.zeros(2, 3), torch.zeros(2, 3), torch.zeros(2, 1), shs=torch.zeros(2, 1, 3))
    assert callable(distCUDA2)


def test_no_cpu_fallback():
    import scorp_b200.diff_gaussian_rasterization as dgr
    from scorp_b200 import _lib
    rs = dgr.GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                           torch.zeros(3), False, False)
    with pytest.raises(_lib.GsrError):
        dgr.GaussianRasterizer(rs)(torch.zeros(2, 3), torch.zeros(2, 3), torch.zeros(2, 1), shs=torch.zeros(2, 1, 3),
                                   scales=torch.ones(2, 3), rotations=torch.ones(2, 4))


def test_product_never_imports_the_oracle():
    for p in (ROOT / "scorp_b200").rglob("*.py"):
        txt = p.read_text()
        assert not re.search(r"^\s*(from|import)\s+oracle\b", txt, flags=re.M), p
    for p in (ROOT / "scorp_b200" / "csrc").glob("*"):
        assert not re.search(r"#\s*include\s*[<\"][^>\"]*oracle", p.read_text()), p      # no oracle header in the kernels


def test_no_cpu_fallback_in_the_rows_next_to_the_boundary():
    """model_space, loss_utils, optim, batched views: CPU tensors raise GsrError, nothing is computed on the host."""
    import scorp_b200.diff_gaussian_rasterization as dgr
    from scorp_b200 import _lib, loss_utils as lu, model_space as ms, optim
    rs = dgr.GaussianRasterizationSettings(8, 8, 1.0, 1.0, torch.zeros(3), 1.0, torch.eye(4), torch.eye(4), 0,
                                           torch.zeros(3), False, False)
    z = torch.zeros
    with pytest.raises(_lib.GsrError):
        ms.rasterize_model(z(2, 3), z(2, 3), z(2, 1, 3), z(2, 0, 3), z(2, 1), z(2, 3), z(2, 4), rs)
    with pytest.raises(_lib.GsrError):
        ms.rasterize_surfel_model(z(2, 3), z(2, 3), z(2, 1, 3), z(2, 0, 3), z(2, 1), z(2, 2), z(2, 4), rs)
    with pytest.raises(_lib.GsrError):
        lu.l1_ssim_loss(z(3, 16, 16), z(3, 16, 16))
    with pytest.raises(_lib.GsrError):
        lu.ssim(z(3, 16, 16), z(3, 16, 16))
    p = torch.nn.Parameter(z(4)); p.grad = z(4)
    with pytest.raises(_lib.GsrError):
        optim.Adam([p], lr=1e-3).step()
    with pytest.raises(_lib.GsrError):
        optim.add_densification_stats(z(4, 1), z(4, 1), z(4), z(4, 3), z(4, dtype=torch.int32))
    with pytest.raises(_lib.GsrError):
        optim.densify_and_prune_tensors(dict(xyz=z(4, 3), f_dc=z(4, 1, 3), f_rest=z(4, 0, 3), opacity=z(4, 1), scaling=z(4, 3),
                                             rotation=torch.ones(4, 4)), {}, z(4, 1), z(4, 1), 0.0002, 0.005, 4.0, 20, 0.01)
    with pytest.raises(_lib.GsrError):
        dgr.GaussianRasterizer(rs).forward_views(torch.eye(4)[None], torch.eye(4)[None], z(1, 3), z(2, 3), z(2, 1),
                                                 shs=z(2, 1, 3), scales=torch.ones(2, 3), rotations=torch.ones(2, 4))


def test_adam_keeps_torch_optimizer_surface():
    """scorp_b200.optim.Adam is a torch.optim.Optimizer with Adam's param_groups / defaults, so
    GaussianModel.update_learning_rate (gaussian_model.py:208-214) and state editing work unchanged."""
    from scorp_b200 import optim
    p = torch.nn.Parameter(torch.zeros(3))
    o = optim.Adam([{"params": [p], "lr": 0.1, "name": "xyz"}], lr=0.0, eps=1e-15)
    assert isinstance(o, torch.optim.Optimizer)
    assert o.param_groups[0]["name"] == "xyz" and o.param_groups[0]["lr"] == 0.1 and o.param_groups[0]["eps"] == 1e-15
    o.param_groups[0]["lr"] = 0.05
    assert o.state_dict()["param_groups"][0]["lr"] == 0.05
    o.step()                                     # no gradient anywhere: nothing to do, no CUDA needed
    with pytest.raises(NotImplementedError):
        optim.Adam([p], amsgrad=True)
