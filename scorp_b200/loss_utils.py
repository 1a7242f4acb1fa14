"""Photometric loss of the training loops on hand-written sm_100a kernels.

Mirror of the reference's gs3dgs/utils/loss_utils.py (same names and argument meaning for the two functions the
training loops call around the rasterizer, train_3dgs.py:106-107):

    l1_loss(network_output, gt)                          loss_utils.py:17-18
    ssim(img1, img2, window_size=11, size_average=True)  loss_utils.py:45-73

plus the fused form the loop actually needs,

    l1_ssim_loss(image, gt, lambda_dssim) == (1 - lambda) * l1_loss(image, gt) + lambda * (1 - ssim(image, gt)),

one forward kernel + one backward kernel (csrc/loss.cu) instead of five grouped conv2d calls and ~20 elementwise
kernels each way. Gradients flow to the first argument only (the rendered image); gt never needs one in the
reference's loops. CUDA tensors only: there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

__all__ = ["l1_loss", "ssim", "l1_ssim_loss", "l1_and_ssim"]


def _prep(img, gt):
    if not img.is_cuda or not gt.is_cuda:
        raise _lib.GsrError("loss kernels need CUDA tensors: there is no CPU path")
    if img.shape != gt.shape:
        raise RuntimeError(f"image shapes differ: {tuple(img.shape)} vs {tuple(gt.shape)}")
    if img.dim() == 3:
        Cn, H, W = img.shape
    elif img.dim() == 4:         # (B, C, H, W): planes are independent, so B*C planes
        Cn, H, W = img.shape[0] * img.shape[1], img.shape[2], img.shape[3]
    else:
        raise RuntimeError("expected (C, H, W) or (B, C, H, W) images")
    f = lambda t: t.detach().to(torch.float32).contiguous()
    return f(img), f(gt), int(Cn), int(H), int(W)


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


class _L1Ssim(torch.autograd.Function):
    """-> out3 = (loss, l1, ssim) as one device tensor; backward accepts a gradient on each of the three."""

    @staticmethod
    def forward(ctx, img, gt, lambda_dssim):
        lib = _lib.lib()
        a, b, Cn, H, W = _prep(img, gt)
        dev = a.device
        nbytes = lib.gsr_l1_ssim_workspace_bytes(Cn, H, W)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        out = torch.empty(3, dtype=torch.float32, device=dev)
        with torch.cuda.device(dev):
            rc = lib.gsr_l1_ssim_forward(Cn, H, W, a.data_ptr(), b.data_ptr(), float(lambda_dssim), out.data_ptr(),
                                         ws.data_ptr(), nbytes, _stream(dev))
        _lib.check(rc, "gsr_l1_ssim_forward")
        ctx.save_for_backward(a, b, ws)
        ctx.dims, ctx.lam, ctx.shape, ctx.dtype = (Cn, H, W), float(lambda_dssim), img.shape, img.dtype
        return out

    @staticmethod
    def backward(ctx, g_out):
        lib = _lib.lib()
        a, b, ws = ctx.saved_tensors
        Cn, H, W = ctx.dims
        dev = a.device
        # d/dimg of g0*loss + g1*l1 + g2*ssim = (g0*(1-lam) + g1) * dl1 + (g2 - g0*lam) * dssim.
        # The common case (only `loss` used) keeps the upstream scalar on the device: no host sync.
        g = g_out.to(torch.float32)
        grad = torch.empty_like(a)
        with torch.cuda.device(dev):
            if ctx.only_loss:
                up = g[0:1].contiguous()
                rc = lib.gsr_l1_ssim_backward(Cn, H, W, a.data_ptr(), b.data_ptr(), up.data_ptr(), 1.0 - ctx.lam,
                                              -ctx.lam, ws.data_ptr(), grad.data_ptr(), _stream(dev))
            else:
                g0, g1, g2 = (float(v) for v in g.tolist())          # general form: three host scalars
                rc = lib.gsr_l1_ssim_backward(Cn, H, W, a.data_ptr(), b.data_ptr(), None, g0 * (1.0 - ctx.lam) + g1,
                                              g2 - g0 * ctx.lam, ws.data_ptr(), grad.data_ptr(), _stream(dev))
        _lib.check(rc, "gsr_l1_ssim_backward")
        return grad.view(ctx.shape).to(ctx.dtype), None, None


class _OnlyLoss(_L1Ssim):
    @staticmethod
    def forward(ctx, img, gt, lam):
        ctx.only_loss = True
        return _L1Ssim.forward(ctx, img, gt, lam)


class _General(_L1Ssim):
    @staticmethod
    def forward(ctx, img, gt, lam):
        ctx.only_loss = False
        return _L1Ssim.forward(ctx, img, gt, lam)


def l1_ssim_loss(image, gt, lambda_dssim=0.2):
    """(1 - lambda) * l1_loss + lambda * (1 - ssim), the loss of train_3dgs.py:107, as a 0-dim tensor."""
    return _OnlyLoss.apply(image, gt, lambda_dssim)[0]


def l1_and_ssim(image, gt):
    """-> (l1, ssim) 0-dim tensors from one pass (both differentiable w.r.t. image)."""
    out = _General.apply(image, gt, 0.0)
    return out[1], out[2]


def l1_loss(network_output, gt):
    """torch.abs(network_output - gt).mean() (loss_utils.py:17-18)."""
    return _General.apply(network_output, gt, 0.0)[1]


def ssim(img1, img2, window_size=11, size_average=True):
    """Mean SSIM with the 11x11 Gaussian window, sigma 1.5 (loss_utils.py:45-73)."""
    if window_size != 11:
        raise NotImplementedError("only the reference's window_size = 11 is built")
    if not size_average:
        raise NotImplementedError("size_average=False is not used on the training path (loss_utils.py:72-73)")
    return _General.apply(img1, img2, 0.0)[2]
