"""TEST INFRASTRUCTURE (checker) for the row DESIGN.md section 10 names next: the maps the 2DGS caller derives from the
rasterizer's 7-channel output, and the two regularisers of the 2DGS training loop. Never imported by scorp_b200/.
No product kernel exists for this row yet; the oracle and its goldens come first.

Follows /root/reference/gs2dgs/gaussian_renderer/__init__.py:128-160 (alpha, world-space normal, median / expected
depth with nan_to_num, distortion, surf_depth, surf_normal), /root/reference/gs2dgs/utils/point_utils.py:9-40
(depths_to_points, depth_to_normal) and /root/reference/train_2dgs.py:142-147 (normal and distortion losses).
float32 with one rounding per torch op where the op is elementwise; the small matrix products and inverses run in
float32 numpy (LAPACK), so values agree with torch to rounding, not bit for bit.
Parity status: depth_to_normal PINNED by tests/golden/surfel_maps.npz = the reference's own function run on CPU
(tools/make_golden_surfel_maps.py); the inline statements of render() and the two losses are not callable without the
CUDA rasterizer, so the golden script replays them in torch and they are pinned only to that replay."""
from __future__ import annotations

import numpy as np

f32 = np.float32


def depths_to_points(world_view_transform, full_proj_transform, W, H, depthmap):
    """point_utils.py:9-26"""
    wvt = np.asarray(world_view_transform, f32); fpt = np.asarray(full_proj_transform, f32)
    c2w = np.linalg.inv(wvt.T).astype(f32)
    ndc2pix = np.array([[W / 2, 0, 0, W / 2], [0, H / 2, 0, H / 2], [0, 0, 0, 1]], f32).T
    projection_matrix = (c2w.T @ fpt).astype(f32)
    intrins = (projection_matrix @ ndc2pix).astype(f32)[:3, :3].T
    gx, gy = np.meshgrid(np.arange(W, dtype=f32), np.arange(H, dtype=f32), indexing="xy")
    points = np.stack([gx, gy, np.ones_like(gx)], axis=-1).reshape(-1, 3)
    rays_d = ((points @ np.linalg.inv(intrins).astype(f32).T).astype(f32) @ c2w[:3, :3].T).astype(f32)
    rays_o = c2w[:3, 3]
    return ((np.asarray(depthmap, f32).reshape(-1, 1) * rays_d).astype(f32) + rays_o).astype(f32)


def depth_to_normal(world_view_transform, full_proj_transform, W, H, depth):
    """point_utils.py:28-40; depth [1, H, W] -> [H, W, 3], border pixels zero"""
    points = depths_to_points(world_view_transform, full_proj_transform, W, H, depth).reshape(H, W, 3)
    out = np.zeros_like(points)
    dx = (points[2:, 1:-1] - points[:-2, 1:-1]).astype(f32)
    dy = (points[1:-1, 2:] - points[1:-1, :-2]).astype(f32)
    n = np.cross(dx, dy).astype(f32)
    nrm = np.maximum(np.sqrt((n * n).sum(-1, keepdims=True, dtype=f32)).astype(f32), f32(1e-12))      # F.normalize
    out[1:-1, 1:-1, :] = (n / nrm).astype(f32)
    return out


def _nan_to_num0(x):
    """torch.nan_to_num(x, 0, 0): nan -> 0, +inf -> 0, -inf stays the most negative float"""
    x = np.array(x, f32, copy=True)
    x[np.isnan(x)] = 0; x[np.isposinf(x)] = 0
    x[np.isneginf(x)] = np.finfo(f32).min
    return x


def surfel_maps(allmap, world_view_transform, full_proj_transform, W, H, depth_ratio):
    """gaussian_renderer/__init__.py:128-160; allmap [7, H, W] -> the five maps render() returns"""
    allmap = np.asarray(allmap, f32)
    wvt = np.asarray(world_view_transform, f32)
    render_alpha = allmap[1:2]
    render_normal = (allmap[2:5].transpose(1, 2, 0) @ wvt[:3, :3].T).astype(f32).transpose(2, 0, 1)
    median = _nan_to_num0(allmap[5:6])
    with np.errstate(divide="ignore", invalid="ignore"):
        expected = _nan_to_num0((allmap[0:1] / render_alpha).astype(f32))
    render_dist = allmap[6:7]
    surf_depth = ((expected * f32(1 - depth_ratio)).astype(f32) + (f32(depth_ratio) * median).astype(f32)).astype(f32)
    surf_normal = depth_to_normal(wvt, full_proj_transform, W, H, surf_depth).transpose(2, 0, 1)
    surf_normal = (surf_normal * render_alpha).astype(f32)
    return dict(render_alpha=render_alpha, render_normal=render_normal, render_dist=render_dist, render_depth=surf_depth,
                surf_normal=surf_normal)


def regularisers(rend_normal, surf_normal, rend_dist, lambda_normal, lambda_dist):
    """train_2dgs.py:145-147 -> (normal_loss, dist_loss) as float64 means of float32 maps"""
    err = f32(1) - (np.asarray(rend_normal, f32) * np.asarray(surf_normal, f32)).sum(0, dtype=f32)
    return lambda_normal * float(err.mean(dtype=np.float64)), lambda_dist * float(np.asarray(rend_dist, f32).mean(dtype=np.float64))
