"""Golden vectors for the 2DGS caller-side maps (the row after SURVEY.md 8(f)): the reference's own
depth_to_normal (/root/reference/gs2dgs/utils/point_utils.py, imported unmodified; matplotlib stubbed, `.cuda()` and
device="cuda" made no-ops for the run) on CPU, plus a torch replay of the inline statements of
gs2dgs/gaussian_renderer/__init__.py:128-160 and train_2dgs.py:142-147, which cannot be called without the CUDA
rasterizer. Writes tests/golden/surfel_maps.npz.   python tools/make_golden_surfel_maps.py"""
import importlib
import math
import sys
import types
from pathlib import Path

import numpy as np
import torch

sys.dont_write_bytecode = True
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from scorp_b200 import synth  # noqa: E402  (camera matrices of the bench scenes)


def _reference():
    sys.path.insert(0, "/root/reference")
    for name in ("matplotlib", "matplotlib.pyplot"):
        sys.modules.setdefault(name, types.ModuleType(name))
    return importlib.import_module("gs2dgs.utils.point_utils")


class _cpu:
    def __enter__(self):
        self.cuda, self.arange = torch.Tensor.cuda, torch.arange
        torch.Tensor.cuda = lambda t, *a, **k: t
        ar = self.arange
        torch.arange = lambda *a, **k: ar(*a, **{x: y for x, y in k.items() if x != "device"})

    def __exit__(self, *e):
        torch.Tensor.cuda, torch.arange = self.cuda, self.arange


def case(W, H, seed, depth_ratio, view):
    pu = _reference()
    g = torch.Generator().manual_seed(seed)
    sc = synth.make_scene("C4", P=16, W=W, H=H, n_views=view + 1)
    cam = sc.cameras[view]
    alpha = torch.rand(1, H, W, generator=g)
    alpha[0, ::7, ::5] = 0.0                                             # empty pixels: 0/0 in the expected depth
    depth = (2.0 + 6.0 * torch.rand(1, H, W, generator=g)) * alpha
    normal = torch.nn.functional.normalize(torch.randn(3, H, W, generator=g), dim=0) * alpha
    median = 2.0 + 6.0 * torch.rand(1, H, W, generator=g)
    median[0, 3::11, 2::9] = float("nan")
    dist = torch.rand(1, H, W, generator=g) * 1e-2
    allmap = torch.cat([depth, alpha, normal, median, dist], 0)
    view_ns = types.SimpleNamespace(world_view_transform=cam.viewmatrix, full_proj_transform=cam.projmatrix, resolution=(W, H))
    # gaussian_renderer/__init__.py:128-160, replayed
    render_alpha = allmap[1:2]
    render_normal = (allmap[2:5].permute(1, 2, 0) @ (view_ns.world_view_transform[:3, :3].T)).permute(2, 0, 1)
    med = torch.nan_to_num(allmap[5:6], 0, 0)
    exp = torch.nan_to_num(allmap[0:1] / render_alpha, 0, 0)
    surf_depth = exp * (1 - depth_ratio) + depth_ratio * med
    with _cpu():
        sn = pu.depth_to_normal(view_ns, surf_depth)                     # the reference's own function
        pts = pu.depths_to_points(view_ns, surf_depth)
    surf_normal = sn.permute(2, 0, 1) * render_alpha.detach()
    # train_2dgs.py:145-147
    lam_n, lam_d = 0.05, 1000.0
    normal_loss = lam_n * (1 - (render_normal * surf_normal).sum(dim=0))[None].mean()
    dist_loss = lam_d * allmap[6:7].mean()
    return dict(allmap=allmap.numpy(), wvt=cam.viewmatrix.numpy(), fpt=cam.projmatrix.numpy(), WH=np.int32([W, H]),
                depth_ratio=np.float64(depth_ratio), points=pts.numpy(), depth_to_normal=sn.numpy(),
                render_normal=render_normal.numpy(), render_depth=surf_depth.numpy(), surf_normal=surf_normal.numpy(),
                losses=np.float64([float(normal_loss), float(dist_loss)]), lambdas=np.float64([lam_n, lam_d]))


def main():
    cases = {"a": case(96, 64, 1, 0.0, 0), "b": case(77, 53, 2, 1.0, 1), "c": case(40, 30, 3, 0.35, 2)}
    np.savez_compressed(ROOT / "tests" / "golden" / "surfel_maps.npz", cases=np.array(sorted(cases)),
                        **{f"{c}_{k}": v for c, d in cases.items() for k, v in d.items()})
    print("wrote tests/golden/surfel_maps.npz", {c: d["allmap"].shape for c, d in cases.items()})


if __name__ == "__main__":
    main()
