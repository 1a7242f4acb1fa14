"""`_C` level of the 3DGS package: same three entry points, argument order and return tuples as
the reference's pybind11 module (reference submodules/diff-gaussian-rasterization/ext.cpp:15-19,
rasterize_points.h:18-70), implemented by calling libgsr_b200.so through its C ABI.

This file owns what the reference's torch glue owned (rasterize_points.cu:36-229): output
allocation, the three scratch buffers (handed to the library through an allocator callback,
like the reference's resize lambdas at rasterize_points.cu:27-33) and pointer extraction.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _lib
from .._lib import GsrSettings, ptr


_backward_hook = [None]      # optional callable(head): runs on the parameter-gradient head of every backward
_factored = [None]           # multiview.install_factored_exchange


def _f32c(t: torch.Tensor, device) -> torch.Tensor:
    """contiguous fp32 on `device`; 0-element sentinels (CPU or CUDA) pass through."""
    if t is None or t.numel() == 0:
        return t
    if t.dtype != torch.float32:
        t = t.float()
    if t.device != device:
        t = t.to(device)
    return t.contiguous()


class _Arena:
    """Allocator handed to the library: one uint8 tensor per scratch buffer.

    The callback is a closure over the buffer dict, NOT a bound method: `self.fn = ALLOC_FN(self._alloc)` would make
    the object reference itself, and the three scratch tensors (140 MB at 1 M Gaussians) would then stay allocated
    until the cyclic garbage collector runs instead of being released with the autograd graph."""

    def __init__(self, device):
        self.device = device
        self.bufs = {}
        self._errors = []
        bufs, errors = self.bufs, self._errors

        def _alloc(_user, which, nbytes):
            try:
                t = torch.empty(max(int(nbytes), 1), dtype=torch.uint8, device=device)
                bufs[which] = t
                return t.data_ptr()
            except Exception as ex:  # never let an exception cross the C ABI
                errors.append(ex)
                return None

        self.fn = _lib.ALLOC_FN(_alloc)

    @property
    def error(self):
        return self._errors[0] if self._errors else None

    def get(self, which):
        t = self.bufs.get(which)
        return t if t is not None else torch.empty(0, dtype=torch.uint8, device=self.device)


def _arm_early(lib, fx, g_drgb, campos, P, rebuild_args=None):
    """Factored multi-GPU exchange: ask the next backward call to hand out the clamped colour gradient right after
    its blend kernel and start gathering it there (gsr_backward_early_rgb). Returns (callback to keep alive until the
    backward returned, list that receives an exception raised inside it); no reference cycle: the closure does not
    see the callback object."""
    if hasattr(fx, "arm"):             # multiview.PeerExchange: the backward pushes drgb into peer memory itself
        armed = fx.arm(lib, campos, P, g_drgb.device, rebuild_args)     # rebuild_args = (means3D, g_sh, M, degree, head)
        if armed is not None:
            return armed
    errors = []

    def _ready(_user):
        try:
            fx.begin(g_drgb, campos, P)
        except Exception as ex:      # never let an exception cross the C ABI
            errors.append(ex)

    cb = _lib.READY_FN(_ready)
    _lib.check(lib.gsr_backward_early_rgb(ptr(g_drgb), cb, None), "gsr_backward_early_rgb")
    return cb, errors


def _settings(bg, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier, degree,
              prefiltered, debug):
    s = GsrSettings()
    s.image_height, s.image_width = int(H), int(W)
    s.tanfovx, s.tanfovy, s.scale_modifier = float(tan_fovx), float(tan_fovy), float(scale_modifier)
    s.sh_degree, s.prefiltered, s.debug = int(degree), int(bool(prefiltered)), int(bool(debug))
    s.bg, s.viewmatrix, s.projmatrix, s.campos = ptr(bg), ptr(viewmatrix), ptr(projmatrix), ptr(campos)
    return s


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def rasterize_gaussians(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                        cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                        image_width, sh, degree, campos, prefiltered, debug):
    """-> (num_rendered, color, depth, alpha, radii, geomBuffer, binningBuffer, imgBuffer)"""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")   # rasterize_points.cu:57-59
    if not means3D.is_cuda:
        raise _lib.GsrError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
    lib = _lib.lib()
    dev = means3D.device
    P, H, W = means3D.size(0), int(image_height), int(image_width)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); opacity = _f32c(opacity, dev)
    scales = _f32c(scales, dev); rotations = _f32c(rotations, dev); cov3D_precomp = _f32c(cov3D_precomp, dev)
    sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)

    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    out_alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _Arena(dev)
    s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                  degree, prefiltered, debug)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        rc = lib.gsr_rasterize_forward(
            C.byref(s), P, M, ptr(means3D), ptr(sh), ptr(colors), ptr(opacity), ptr(scales), ptr(rotations),
            ptr(cov3D_precomp), out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(),
            ptr(radii), arena.fn, None, C.byref(n), _stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_rasterize_forward")
    return (n.value, out_color, out_depth, out_alpha, radii,
            arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING), arena.get(_lib.GSR_BUF_IMAGE))


def rasterize_gaussians_nosync(background, means3D, colors, opacity, scales, rotations, scale_modifier,
                               cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy, image_height,
                               image_width, sh, degree, campos, prefiltered, debug, capacity, status):
    """rasterize_gaussians without the host round trip (include/gsr_b200.h gsr_rasterize_forward_nosync; no reference
    counterpart). `capacity`: instances the binning buffer holds; `status`: pinned int32[4] that receives
    {num_rendered, longest list, overflow, -} asynchronously (scorp_b200/capacity.py). Returns the same tuple with
    num_rendered replaced by `capacity` (what the backward must be given as R)."""
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
    scales = _f32c(scales, dev); rotations = _f32c(rotations, dev); cov3D_precomp = _f32c(cov3D_precomp, dev)
    sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)
    out_color = torch.empty((3, H, W), dtype=torch.float32, device=dev)
    out_depth = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    out_alpha = torch.empty((1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((P,), dtype=torch.int32, device=dev)
    arena = _Arena(dev)
    s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                  degree, prefiltered, debug)
    with torch.cuda.device(dev):
        rc = lib.gsr_rasterize_forward_nosync(
            C.byref(s), P, M, ptr(means3D), ptr(sh), ptr(colors), ptr(opacity), ptr(scales), ptr(rotations),
            ptr(cov3D_precomp), out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(),
            ptr(radii), arena.fn, None, int(capacity), status.data_ptr(), _stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_rasterize_forward_nosync")
    return (int(capacity), out_color, out_depth, out_alpha, radii,
            arena.get(_lib.GSR_BUF_GEOM), arena.get(_lib.GSR_BUF_BINNING), arena.get(_lib.GSR_BUF_IMAGE))


def rasterize_gaussians_views(background, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D_precomp,
                              viewmatrices, projmatrices, tan_fovx, tan_fovy, image_height, image_width, sh, degree,
                              camposes, prefiltered, debug):
    """Batched forward (no reference counterpart; include/gsr_b200.h gsr_rasterize_forward_views): same arguments as
    rasterize_gaussians with viewmatrices[V,4,4], projmatrices[V,4,4], camposes[V,3].
    -> (num_rendered, color[V,3,H,W], depth[V,1,H,W], alpha[V,1,H,W], radii[V,P]); forward only."""
    if means3D.dim() != 2 or means3D.size(1) != 3:
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    if not means3D.is_cuda:
        raise _lib.GsrError("means3D must be a CUDA tensor: this rasterizer has no CPU path")
    if viewmatrices.dim() != 3 or viewmatrices.shape[1:] != (4, 4) or projmatrices.shape != viewmatrices.shape \
            or camposes.shape != (viewmatrices.size(0), 3):
        raise RuntimeError("expected viewmatrices[V,4,4], projmatrices[V,4,4], camposes[V,3]")
    lib = _lib.lib()
    dev = means3D.device
    V, P, H, W = viewmatrices.size(0), means3D.size(0), int(image_height), int(image_width)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); opacity = _f32c(opacity, dev)
    scales = _f32c(scales, dev); rotations = _f32c(rotations, dev); cov3D_precomp = _f32c(cov3D_precomp, dev)
    sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrices = _f32c(viewmatrices, dev)
    projmatrices = _f32c(projmatrices, dev); camposes = _f32c(camposes, dev)
    out_color = torch.empty((V, 3, H, W), dtype=torch.float32, device=dev)
    out_depth = torch.empty((V, 1, H, W), dtype=torch.float32, device=dev)
    out_alpha = torch.empty((V, 1, H, W), dtype=torch.float32, device=dev)
    radii = torch.empty((V, P), dtype=torch.int32, device=dev)
    if V == 0:
        return 0, out_color, out_depth, out_alpha, radii
    arena = _Arena(dev)
    s = _settings(background, viewmatrices, projmatrices, camposes, tan_fovx, tan_fovy, H, W, scale_modifier, degree,
                  prefiltered, debug)
    n = C.c_int(0)
    with torch.cuda.device(dev):
        rc = lib.gsr_rasterize_forward_views(
            C.byref(s), V, P, M, ptr(means3D), ptr(sh), ptr(colors), ptr(opacity), ptr(scales), ptr(rotations),
            ptr(cov3D_precomp), out_color.data_ptr(), out_depth.data_ptr(), out_alpha.data_ptr(), ptr(radii),
            arena.fn, None, C.byref(n), _stream(dev))
    if arena.error is not None:
        raise arena.error
    _lib.check(rc, "gsr_rasterize_forward_views")
    return n.value, out_color, out_depth, out_alpha, radii


def rasterize_gaussians_backward(background, means3D, radii, colors, scales, rotations, scale_modifier,
                                 cov3D_precomp, viewmatrix, projmatrix, tan_fovx, tan_fovy,
                                 dL_dout_color, dL_dout_depth, dL_dout_alpha, sh, degree, campos,
                                 geomBuffer, R, binningBuffer, imageBuffer, alpha, debug):
    """-> (dL_dmeans2D, dL_dcolors, dL_dopacity, dL_dmeans3D, dL_dcov3D, dL_dsh, dL_dscales, dL_drotations)"""
    lib = _lib.lib()
    dev = means3D.device
    P = means3D.size(0)
    H, W = dL_dout_color.size(1), dL_dout_color.size(2)
    M = sh.size(1) if sh is not None and sh.numel() != 0 else 0
    means3D = _f32c(means3D, dev); colors = _f32c(colors, dev); scales = _f32c(scales, dev)
    rotations = _f32c(rotations, dev); cov3D_precomp = _f32c(cov3D_precomp, dev); sh = _f32c(sh, dev)
    background = _f32c(background, dev); viewmatrix = _f32c(viewmatrix, dev)
    projmatrix = _f32c(projmatrix, dev); campos = _f32c(campos, dev)
    dL_dout_color = _f32c(dL_dout_color, dev); dL_dout_depth = _f32c(dL_dout_depth, dev)
    dL_dout_alpha = _f32c(dL_dout_alpha, dev); alpha = _f32c(alpha, dev)

    # One arena for all gradients (fully written by the library: no zero fill needed). The five
    # parameter gradients come first and are contiguous, so a view-sharded trainer can sum them
    # across ranks with ONE allreduce over arena[:P * PARAM_GRAD_FLOATS(M)].
    # With multiview.install_factored_exchange the SH gradient is NOT produced by the backward (dL_dsh = NULL): the
    # arena head is [means3D, opacity, scales, rotations] (all-reduced), followed by the clamped colour gradient that
    # is all-gathered, and dL_dsh is rebuilt from all ranks' colour gradients (csrc/sh_exchange.cu).
    fx_installed = _factored[0]
    fx = fx_installed if (M > 0 and P != 0) else None
    if fx is not None:
        widths = (3, 1, 3, 4, 3, 3, 3, 6)       # means3D, opacity, scales, rotations | drgb | means2D, colors, cov3D
    else:
        widths = (3, 3 * M, 1, 3, 4, 3, 3, 6)   # means3D, sh, opacity, scales, rotations | means2D, colors, cov3D
    flat = torch.empty((P * sum(widths),), dtype=torch.float32, device=dev)
    outs, o = [], 0
    for w in widths:
        outs.append(flat[o:o + P * w]); o += P * w
    if fx is not None:
        g_m3d, g_op, g_sc, g_rot, g_drgb, g_m2d, g_col, g_cov = outs
        g_sh = torch.empty((P * M * 3,), dtype=torch.float32, device=dev)
    else:
        g_m3d, g_sh, g_op, g_sc, g_rot, g_m2d, g_col, g_cov = outs
    if P != 0:
        ws_bytes = lib.gsr_backward_workspace_bytes(P)
        ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
        s = _settings(background, viewmatrix, projmatrix, campos, tan_fovx, tan_fovy, H, W, scale_modifier,
                      degree, False, debug)
        early = _arm_early(lib, fx, g_drgb, campos, P, (means3D, g_sh, M, int(degree), flat[: P * 11])) if fx is not None else None
        with torch.cuda.device(dev):
            rc = lib.gsr_rasterize_backward(
                C.byref(s), P, M, int(R), ptr(means3D), ptr(sh), ptr(colors), ptr(scales), ptr(rotations),
                ptr(cov3D_precomp), ptr(radii), ptr(geomBuffer), ptr(binningBuffer), ptr(imageBuffer),
                ptr(alpha), ptr(dL_dout_color), ptr(dL_dout_depth), ptr(dL_dout_alpha),
                ptr(g_m2d), ptr(g_col), ptr(g_op), ptr(g_m3d), ptr(g_cov), None if fx is not None else ptr(g_sh),
                ptr(g_sc), ptr(g_rot), ws.data_ptr(), ws_bytes, _stream(dev))
            _lib.check(rc, "gsr_rasterize_backward")
        if fx is not None:
            if early[1]:
                raise early[1][0]
            fx.finish(flat[: P * 11], means3D, g_sh, P, M, int(degree))
        elif fx_installed is not None:
            # precomputed colours (M == 0): there is no SH gradient to factor, but install_factored_exchange promises
            # gradients summed over ranks from EVERY backward, so the head is reduced with one all_reduce in its group
            from .. import multiview as _mv
            _mv.reduce_without_factoring(fx_installed, flat[: P * param_grad_floats(M)])
        elif _backward_hook[0] is not None:     # e.g. multiview.install_backward_allreduce: reduce before autograd copies
            _backward_hook[0](flat[: P * param_grad_floats(M)])
    return (g_m2d.view(P, 3), g_col.view(P, 3), g_op.view(P, 1), g_m3d.view(P, 3), g_cov.view(P, 6),
            g_sh.view(P, M, 3), g_sc.view(P, 3), g_rot.view(P, 4))


def param_grad_floats(M: int) -> int:
    """floats per Gaussian at the head of the gradient arena: means3D, sh, opacity, scales, rotations."""
    return 3 + 3 * M + 1 + 3 + 4


def grad_arena(grad_means3D: torch.Tensor) -> torch.Tensor:
    """The flat arena a backward result lives in (its first P*param_grad_floats(M) floats are the
    parameter gradients)."""
    base = grad_means3D._base
    while base is not None and base._base is not None:
        base = base._base
    return base


def mark_visible(means3D, viewmatrix, projmatrix):
    """bool[P]: view-space z > 0.2 (rasterize_points.cu:210-229)."""
    lib = _lib.lib()
    if not means3D.is_cuda:
        raise _lib.GsrError("means3D must be a CUDA tensor")
    dev = means3D.device
    P = means3D.size(0)
    present = torch.zeros((P,), dtype=torch.bool, device=dev)
    if P != 0:
        means3D = _f32c(means3D, dev); viewmatrix = _f32c(viewmatrix, dev); projmatrix = _f32c(projmatrix, dev)
        with torch.cuda.device(dev):
            rc = lib.gsr_mark_visible(P, ptr(means3D), ptr(viewmatrix), ptr(projmatrix), present.data_ptr(),
                                      _stream(dev))
        _lib.check(rc, "gsr_mark_visible")
    return present
