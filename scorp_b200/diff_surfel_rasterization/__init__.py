"""Drop-in `diff_surfel_rasterization` (2DGS) backed by hand-written sm_100a CUDA (libgsr_b200.so).

Same public surface as the reference package
(reference submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py:158-222):
GaussianRasterizationSettings, GaussianRasterizer.forward(...) -> (color[3,H,W], radii[P], allmap[7,H,W])
with allmap channels depth 0, alpha 1, normal 2-4, median depth 5, distortion 6
(cuda_rasterizer/auxiliary.h:23-27); scales are (P,2); cov3D_precomp means a (P,9) transMat.
"""
from __future__ import annotations

import torch
import torch.nn as nn

from ..diff_gaussian_rasterization import GaussianRasterizationSettings, _call
from . import _C
from .. import capacity as _capacity

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        # scorp_b200.capacity.enable(): no host round trip once the instance count of this shape is known
        key = ("2dgs", int(means3D.shape[0]), int(rs.image_width), int(rs.image_height))
        cap = _capacity.capacity_for(key) if (_capacity.enabled() and not rs.debug and means3D.shape[0] > 0) else None
        if cap is not None:
            slot = _capacity.status_slot()
            (num_rendered, color, others, radii, geom, binning, img) = _C.rasterize_gaussians_nosync(*args, cap, slot)
            _capacity.note(key, slot, cap)
        else:
            (num_rendered, color, others, radii, geom, binning, img) = _call(
                _C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
            if _capacity.enabled():
                _capacity.observe(key, num_rendered)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img)
        ctx.mark_non_differentiable(radii)
        return color, radii, others

    @staticmethod
    def backward(ctx, grad_out_color, grad_radii, grad_others):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img) = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        if grad_out_color is None:
            grad_out_color = torch.zeros((3, H, W), dtype=torch.float32, device=means3D.device)
        if grad_others is None:
            grad_others = torch.zeros((_C.NUM_OTHERS, H, W), dtype=torch.float32, device=means3D.device)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_out_color, grad_others, sh,
                rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, rs.debug)
        (g_means2D, g_colors, g_opacities, g_means3D, g_transMat, g_sh, g_scales, g_rotations) = _call(
            _C.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_transMat, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None and rotations is not None
        any_sr = scales is not None or rotations is not None
        if (not have_sr and cov3D_precomp is None) or (any_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.empty(0, device=means3D.device)     # the reference passes CUDA empties here (:199-209)
        return rasterize_gaussians(
            means3D, means2D,
            empty if shs is None else shs,
            empty if colors_precomp is None else colors_precomp,
            opacities,
            empty if scales is None else scales,
            empty if rotations is None else rotations,
            empty if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
