"""`_C` level of the 2DGS surfel package: same entry points, argument order and return tuples as the
reference's pybind11 module (reference submodules/diff-surfel-rasterization/ext.cpp,
rasterize_points.cu:39-134 forward, :136-232 backward, :234-253 markVisible), implemented by calling
libgsr_b200.so through its C ABI. scales are (P,2); `transMat_precomp` (P,9) takes cov3D_precomp's
place; forward returns one 7-channel map (depth, alpha, normal xyz, median depth, distortion)."""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib
from .._lib import ptr
from ..diff_gaussian_rasterization._C import _Arena, _arm_early, _f32c, _settings, _stream, mark_visible  # noqa: F401

NUM_OTHERS = 7
_backward_hook = [None]      # optional callable(head): runs on the parameter-gradient head of every backward
_factored = [None]           # multiview.install_factored_exchange


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, prefiltered, debug):
    """-> (num_rendered, color, others[7,H,W], radii, geomBuffer, binningBuffer, imgBuffer)"""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise _lib.GsrError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
    lib = _lib.lib()
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); opacity = _f32c(opacity, dev)
    scales = _f32c(scales, dev); rotations = _f32c(rotations, dev); transMat_precomp = _f32c(transMat_precomp, dev)
    sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_others = torch.empty((NUM_OTHERS, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _Arena(dev)
    s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                  degree, prefiltered, debug)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        rc = lib.gsr_surfel_forward(
            C.byref(s), P, M, ptr(means3D), ptr(sh), ptr(colors), ptr(opacity), ptr(scales), ptr(rotations),
            ptr(transMat_precomp), out_color.data_ptr(), out_others.data_ptr(), ptr(radii),
            arena.fn, None, C.byref(n), _stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_surfel_forward")
    return (n.value, out_color, out_others, radii,
            arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING), arena.get(_lib.GSR_BUF_IMAGE))


def rasterize_gaussians_nosync(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                               transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                               image_width, sh, degree, campos, prefiltered, debug, capacity, status):
    """rasterize_gaussians without the host round trip (gsr_surfel_forward_nosync; see the 3DGS `_C` level and
    scorp_b200/capacity.py). Returns the same tuple with num_rendered replaced by `capacity`."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise _lib.GsrError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
    if not (status.is_pinned() and status.dtype == torch.int32 and status.numel() >= 4 and status.is_contiguous()):
        raise _lib.GsrError("status must be a pinned, contiguous int32 tensor of 4 elements")
    lib = _lib.lib()
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); opacity = _f32c(opacity, dev)
    scales = _f32c(scales, dev); rotations = _f32c(rotations, dev); transMat_precomp = _f32c(transMat_precomp, dev)
    sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_others = torch.empty((NUM_OTHERS, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _Arena(dev)
    s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                  degree, prefiltered, debug)
    with torch.cuda.device(dev):
        rc = lib.gsr_surfel_forward_nosync(
            C.byref(s), P, M, ptr(means3D), ptr(sh), ptr(colors), ptr(opacity), ptr(scales), ptr(rotations),
            ptr(transMat_precomp), out_color.data_ptr(), out_others.data_ptr(), ptr(radii),
            arena.fn, None, int(capacity), status.data_ptr(), _stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_surfel_forward_nosync")
    return (int(capacity), out_color, out_others, radii,
            arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING), arena.get(_lib.GSR_BUF_IMAGE))


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 transMat_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_others, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, debug):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dtransMat, dL_dsh, dL_dscales[P,2], dL_drotations)"""
    lib = _lib.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); scales = _f32c(scales, dev)
    rotations = _f32c(rotations, dev); transMat_precomp = _f32c(transMat_precomp, dev); sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)
    dL_dout_color = _f32c(dL_dout_color, dev); dL_dout_others = _f32c(dL_dout_others, dev)
    # parameter gradients first and contiguous (one allreduce), then the rest; with
    # multiview.install_factored_exchange the SH rows are rebuilt from all ranks' colour gradients instead
    # (see diff_gaussian_rasterization/_C.py and csrc/sh_exchange.cu)
    fx_installed = _factored[0]
    fx = fx_installed if (M > 0 and P != 0) else None
    if fx is not None:
        widths = (3, 1, 2, 4, 3, 3, 3, 9)       # means3D, opacity, scales, rotations | drgb | means2D, colors, transMat
    else:
        widths = (3, 3 * M, 1, 2, 4, 3, 3, 9)   # means3D, sh, opacity, scales, rotations | means2D, colors, transMat
    flat = torch.empty((P * sum(widths),), dtype=torch.float32, device=dev)
    outs, o = [], 0
    for w in widths:
        outs.append(flat[o:o + P * w]); o += P * w
    if fx is not None:
        g_m3d, g_op, g_sc, g_rot, g_drgb, g_m2d, g_col, g_tm = outs
        g_sh = torch.empty((P * M * 3,), dtype=torch.float32, device=dev)
    else:
        g_m3d, g_sh, g_op, g_sc, g_rot, g_m2d, g_col, g_tm = outs
    if P != 0:
        ws_bytes = lib.gsr_surfel_backward_workspace_bytes(P)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                      degree, False, debug)
        early = _arm_early(lib, fx, g_drgb, campos, P, (means3D, g_sh, M, int(degree), flat[: P * 10])) if fx is not None else None
        with torch.cuda.device(dev):
            rc = lib.gsr_surfel_backward(
                C.byref(s), P, M, int(R), ptr(means3D), ptr(sh), ptr(colors), ptr(scales), ptr(rotations),
                ptr(transMat_precomp), ptr(radii), ptr(geomBuffer), ptr(binningBuffer), ptr(imageBuffer),
                ptr(dL_dout_color), ptr(dL_dout_others),
                ptr(g_m2d), ptr(g_col), ptr(g_op), ptr(g_m3d), ptr(g_tm), None if fx is not None else ptr(g_sh),
                ptr(g_sc), ptr(g_rot), ws.data_ptr(), ws_bytes, _stream(dev))
            _lib.check(rc, "gsr_surfel_backward")
        if fx is not None:
            if early[1]:
                raise early[1][0]
            fx.finish(flat[: P * 10], means3D, g_sh, P, M, int(degree))
        elif fx_installed is not None:
            # precomputed colours (M == 0): there is no SH gradient to factor, but install_factored_exchange promises
            # gradients summed over ranks from EVERY backward, so the head is reduced with one all_reduce in its group
            from .. import multiview as _mv
            _mv.reduce_without_factoring(fx_installed, flat[: P * param_grad_floats(M)])
        elif _backward_hook[0] is not None:     # e.g. multiview.install_backward_allreduce: reduce before autograd copies
            _backward_hook[0](flat[: P * param_grad_floats(M)])
    return (g_m2d.view(P, 3), g_col.view(P, 3), g_op.view(P, 1), g_m3d.view(P, 3), g_tm.view(P, 9),
            g_sh.view(P, M, 3), g_sc.view(P, 2), g_rot.view(P, 4))


def param_grad_floats(M: int) -> int:
    return 3 + 3 * M + 1 + 2 + 4
