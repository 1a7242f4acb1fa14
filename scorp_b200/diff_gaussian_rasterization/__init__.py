"""Drop-in `diff_gaussian_rasterization` backed by hand-written sm_100a CUDA (libgsr_b200.so).

Same public surface as the reference package
(reference submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py):
  GaussianRasterizationSettings  (:160-172)   12-field NamedTuple
  GaussianRasterizer             (:174-223)   nn.Module with forward(...) and markVisible(...)
  rasterize_gaussians            (:21-42)     functional form
forward returns (color[3,H,W], radii[P] int32, depth[1,H,W], alpha[1,H,W]); gradients come back
in input order; the dummy `means2D` input receives the screen-space positional gradient.
Backward only reads what forward saved, so it may run several times per forward
(reference utils/mask.py:52-89 does that with retain_graph=True).
"""
from __future__ import annotations

from typing import NamedTuple

import torch
import torch.nn as nn

from . import _C
from .. import capacity as _capacity

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _snapshot(args):
    """CPU copies of an argument tuple, taken BEFORE the call (reference __init__.py:17-19)."""
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _call(fn, args, debug, dump_name, phase):
    """Run a `_C` entry point; with debug on, dump the inputs if it raises (reference :83-90,135-142)."""
    if not debug:
        return fn(*args)
    saved = _snapshot(args)
    try:
        return fn(*args)
    except Exception:
        torch.save(saved, dump_name)
        print(f"\nAn error occured in {phase}. Please forward {dump_name} for debugging.")
        raise


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                raster_settings):
        rs = raster_settings
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, rs.image_height,
                rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        # scorp_b200.capacity.enable(): no host round trip once the instance count of this shape is known
        key = ("3dgs", int(means3D.shape[0]), int(rs.image_width), int(rs.image_height))
        cap = _capacity.capacity_for(key) if (_capacity.enabled() and not rs.debug and means3D.shape[0] > 0) else None
        if cap is not None:
            slot = _capacity.status_slot()
            (num_rendered, color, depth, alpha, radii, geom, binning, img) = _C.rasterize_gaussians_nosync(*args, cap, slot)
            _capacity.note(key, slot, cap)
        else:
            (num_rendered, color, depth, alpha, radii, geom, binning, img) = _call(
                _C.rasterize_gaussians, args, rs.debug, "snapshot_fw.dump", "forward")
            if _capacity.enabled():
                _capacity.observe(key, num_rendered)
        ctx.raster_settings = rs
        ctx.num_rendered = num_rendered
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img, alpha)
        ctx.mark_non_differentiable(radii)
        return color, radii, depth, alpha

    @staticmethod
    def backward(ctx, grad_color, grad_radii, grad_depth, grad_alpha):
        rs = ctx.raster_settings
        (colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
         geom, binning, img, alpha) = ctx.saved_tensors
        H, W = rs.image_height, rs.image_width
        # autograd hands None for outputs that did not take part in the loss
        if grad_color is None:
            grad_color = torch.zeros((3, H, W), dtype=torch.float32, device=means3D.device)
        if grad_depth is None:
            grad_depth = torch.zeros((1, H, W), dtype=torch.float32, device=means3D.device)
        if grad_alpha is None:
            grad_alpha = torch.zeros((1, H, W), dtype=torch.float32, device=means3D.device)
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier, cov3Ds_precomp,
                rs.viewmatrix, rs.projmatrix, rs.tanfovx, rs.tanfovy, grad_color, grad_depth, grad_alpha,
                sh, rs.sh_degree, rs.campos, geom, ctx.num_rendered, binning, img, alpha, rs.debug)
        (g_means2D, g_colors, g_opacities, g_means3D, g_cov3D, g_sh, g_scales, g_rotations) = _call(
            _C.rasterize_gaussians_backward, args, rs.debug, "snapshot_bw.dump", "backward")
        return (g_means3D, g_means2D, g_sh, g_colors, g_opacities, g_scales, g_rotations, g_cov3D, None)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations, cov3Ds_precomp,
                        raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                                     cov3Ds_precomp, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: frustum test of the camera in raster_settings (view-space z > 0.2)."""
        with torch.no_grad():
            rs = self.raster_settings
            return _C.mark_visible(positions, rs.viewmatrix, rs.projmatrix)

    def forward_views(self, viewmatrices, projmatrices, camposes, means3D, opacities, shs=None, colors_precomp=None,
                      scales=None, rotations=None, cov3D_precomp=None):
        """Extension (no reference counterpart): render the same Gaussians from V cameras of this rasterizer's
        resolution / field of view in ONE call -- the pose sweeps of align_3dgs_clpe_9dof.py:146-160 and
        utils/mask.py:120-122, which loop over forward(). viewmatrices[V,4,4], projmatrices[V,4,4], camposes[V,3]
        replace the three camera fields of raster_settings. Forward only (runs under no_grad).
        -> (color[V,3,H,W], radii[V,P], depth[V,1,H,W], alpha[V,1,H,W]); slice v equals forward() with camera v."""
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None and rotations is not None
        any_sr = scales is not None or rotations is not None
        if (not have_sr and cov3D_precomp is None) or (any_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        rs = self.raster_settings
        empty = torch.Tensor([])
        with torch.no_grad():
            _, color, depth, alpha, radii = _C.rasterize_gaussians_views(
                rs.bg, means3D, empty if colors_precomp is None else colors_precomp, opacities,
                empty if scales is None else scales, empty if rotations is None else rotations, rs.scale_modifier,
                empty if cov3D_precomp is None else cov3D_precomp, viewmatrices, projmatrices, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, empty if shs is None else shs, rs.sh_degree, camposes,
                rs.prefiltered, rs.debug)
        return color, radii, depth, alpha

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None):
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        have_sr = scales is not None and rotations is not None
        any_sr = scales is not None or rotations is not None
        if (not have_sr and cov3D_precomp is None) or (any_sr and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        empty = torch.Tensor([])      # the reference's "absent" sentinel: a 0-element tensor
        return rasterize_gaussians(
            means3D, means2D,
            empty if shs is None else shs,
            empty if colors_precomp is None else colors_precomp,
            opacities,
            empty if scales is None else scales,
            empty if rotations is None else rotations,
            empty if cov3D_precomp is None else cov3D_precomp,
            self.raster_settings)
