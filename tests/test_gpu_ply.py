"""save_ply / load_ply on the GPU (last item): the packed vertex rows and the file bytes against
the numpy oracle's restatement of GaussianModel.save_ply (gaussian_model.py:234-251), load_ply against its restatement
of :287-327 on files with the reference's layout, a shuffled property order and mixed property types."""
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import ply_oracle as PO

pytestmark = pytest.mark.gpu


def _model(P, R, S, seed=0):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g).cuda()
    return dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, R, 3), opacity=r(P, 1), scaling=r(P, S), rotation=r(P, 4))


def _np(m):
    return {k: v.detach().cpu().numpy() for k, v in m.items()}


@pytest.mark.parametrize("P,R,S,deg", [(10_007, 15, 3, 3), (4_001, 0, 3, 0), (6_113, 8, 2, 2), (1, 3, 3, 1), (0, 15, 3, 3)])
def test_save_matches_oracle_bytes_and_round_trips(tmp_path, P, R, S, deg):
    from scorp_b200 import ply_io
    m = _model(P, R, S, seed=P % 13)
    path = str(tmp_path / "sub" / "point_cloud.ply")            # save_ply creates the directory (mkdir_p, :235)
    ply_io.save_ply(m, path)
    names, rows = PO.vertex_rows(_np(m))
    PO.write_ply(str(tmp_path / "oracle.ply"), names, rows)
    assert open(path, "rb").read() == open(tmp_path / "oracle.ply", "rb").read()
    back = ply_io.load_ply(path, max_sh_degree=deg)
    ref = PO.model_from_columns(names, rows, deg)
    for k in m:
        assert back[k].is_cuda and back[k].shape == m[k].shape, k
        assert torch.equal(back[k], m[k]) and np.array_equal(back[k].cpu().numpy(), ref[k]), k


def test_load_any_property_order_and_types(tmp_path):
    from scorp_b200 import ply_io
    m = _np(_model(5_003, 15, 3, seed=2))
    names, rows = PO.vertex_rows(m)
    rows[:, 3:6] = np.arange(3)                                 # normals present in foreign files are ignored
    perm = list(np.random.default_rng(3).permutation(len(names)))
    PO.write_ply(str(tmp_path / "perm.ply"), [names[i] for i in perm], rows[:, perm], comments=("exported elsewhere",))
    PO.write_ply(str(tmp_path / "mixed.ply"), names, rows, types={"opacity": "f8", "nx": "u1", "scale_1": "f8"})
    PO.write_ply(str(tmp_path / "be.ply"), names, rows, byte_order=">")
    ref = PO.model_from_columns(names, rows, 3)
    for f in ("perm", "mixed", "be"):
        got = ply_io.load_ply(str(tmp_path / f"{f}.ply"), max_sh_degree=3)
        for k in ref:
            assert np.array_equal(got[k].cpu().numpy(), ref[k]), (f, k)
    with pytest.raises(AssertionError):                          # the reference's assert on the SH degree (:304)
        ply_io.load_ply(str(tmp_path / "perm.ply"), max_sh_degree=2)
    keep = [i for i, n in enumerate(names) if n != "opacity"]
    PO.write_ply(str(tmp_path / "noop.ply"), [names[i] for i in keep], rows[:, keep])
    with pytest.raises(ValueError, match="opacity"):
        ply_io.load_ply(str(tmp_path / "noop.ply"))


def test_model_object_surface(tmp_path):
    """save_ply / load_ply_into on a GaussianModel-shaped object, as scene/__init__.py uses them"""
    from scorp_b200 import ply_io
    m = _model(2_000, 15, 3, seed=6)
    pc = SimpleNamespace(_xyz=torch.nn.Parameter(m["xyz"]), _features_dc=torch.nn.Parameter(m["features_dc"]),
                         _features_rest=torch.nn.Parameter(m["features_rest"]), _opacity=torch.nn.Parameter(m["opacity"]),
                         _scaling=torch.nn.Parameter(m["scaling"]), _rotation=torch.nn.Parameter(m["rotation"]))
    path = str(tmp_path / "pc.ply")
    ply_io.save_ply(pc, path)
    new = SimpleNamespace(max_sh_degree=3, active_sh_degree=0)
    ply_io.load_ply_into(new, path, spatial_lr_scale=2.5)
    assert new.active_sh_degree == 3 and new.spatial_lr_scale == 2.5
    assert new.max_radii2D.shape == (2_000,) and float(new.max_radii2D.abs().max()) == 0
    for a in ("_xyz", "_features_dc", "_features_rest", "_opacity", "_scaling", "_rotation"):
        p = getattr(new, a)
        assert isinstance(p, torch.nn.Parameter) and p.requires_grad and torch.equal(p.detach(), getattr(pc, a).detach()), a
