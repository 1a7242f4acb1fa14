"""Optimizer step and densification statistics on hand-written sm_100a kernels.

`Adam` is a drop-in for the optimizer the reference builds in GaussianModel.training_setup
(gs3dgs/scene/gaussian_model.py:192-206): `torch.optim.Adam(l, lr=0.0, eps=1e-15)` with one parameter group per
stored tensor. Same constructor arguments, same `param_groups` (so update_learning_rate's `param_group['lr'] = lr`
works), same `state[p]` keys (`step`, `exp_avg`, `exp_avg_sq`: the densification code edits them in place,
gaussian_model.py cat_tensors_to_optimizer / _prune_optimizer / replace_tensor_to_optimizer) and the same
state_dict layout, so checkpoints interchange with torch.optim.Adam. `step()` updates every group with ONE kernel
launch (csrc/optim.cu) and is bit-identical to torch's default (foreach) implementation.

`add_densification_stats(model_or_tensors, ...)` is train_3dgs.py:180-181 + gaussian_model.py:603-605 in one launch
without the host synchronisation of boolean-mask indexing. CUDA tensors only: there is no CPU path.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

__all__ = ["Adam", "add_densification_stats"]


class Adam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=0, amsgrad=False):
        if weight_decay != 0 or amsgrad:
            raise NotImplementedError("weight_decay / amsgrad are off in the reference's training setup and not built")
        if not 0.0 <= lr or not 0.0 <= eps or not (0.0 <= betas[0] < 1.0) or not (0.0 <= betas[1] < 1.0):
            raise ValueError("invalid Adam hyper-parameter")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=0, amsgrad=False))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        lib = _lib.lib()
        # groups that share (betas, eps, step) go into one launch; the reference's six groups always do
        batches = {}
        for group in self.param_groups:
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise _lib.GsrError("scorp_b200.optim.Adam needs CUDA parameters: there is no CPU path")
                if p.grad.is_sparse or p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("scorp_b200.optim.Adam handles dense contiguous fp32 parameters")
                st = self.state[p]
                if len(st) == 0:
                    st["step"] = torch.tensor(0.0, dtype=torch.float32)      # host scalar, as torch keeps it
                    st["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    st["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                st["step"] += 1
                k = int(st["step"].item()) if isinstance(st["step"], torch.Tensor) else int(st["step"])
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                key = (p.device, group["betas"], group["eps"], k)
                batches.setdefault(key, []).append((p, g, st["exp_avg"], st["exp_avg_sq"], float(group["lr"])))
        for (dev, betas, eps, k), items in batches.items():
            for i in range(0, len(items), 8):
                chunk = items[i:i + 8]
                arr = (_lib.GsrAdamGroup * len(chunk))()
                for a, (p, g, m, v, lr) in zip(arr, chunk):
                    a.param, a.grad, a.exp_avg, a.exp_avg_sq = p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr()
                    a.n, a.lr = p.numel(), lr
                with torch.cuda.device(dev):
                    rc = lib.gsr_adam_step(len(chunk), arr, float(betas[0]), float(betas[1]), float(eps), k,
                                           C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
                _lib.check(rc, "gsr_adam_step")
        return loss


def add_densification_stats(xyz_gradient_accum, denom, max_radii2D, viewspace_grad, radii):
    """In place, for the Gaussians with radii > 0:
        max_radii2D = max(max_radii2D, radii)                        (train_3dgs.py:180; pass None to skip)
        xyz_gradient_accum += norm(viewspace_grad[:, :2], dim=-1)    (gaussian_model.py:604)
        denom += 1                                                   (gaussian_model.py:605)
    viewspace_grad is viewspace_point_tensor.grad ([P, 3])."""
    lib = _lib.lib()
    if not viewspace_grad.is_cuda:
        raise _lib.GsrError("densification statistics need CUDA tensors: there is no CPU path")
    P = radii.numel()
    for t in (xyz_gradient_accum, denom) + ((max_radii2D,) if max_radii2D is not None else ()):
        if t.dtype != torch.float32 or not t.is_contiguous() or t.numel() != P:
            raise RuntimeError("statistics tensors must be contiguous fp32 with one element per Gaussian")
    if radii.dtype != torch.int32 or viewspace_grad.dtype != torch.float32 or viewspace_grad.shape != (P, 3):
        raise RuntimeError("expected radii int32[P] and viewspace gradient fp32[P, 3]")
    g = viewspace_grad if viewspace_grad.is_contiguous() else viewspace_grad.contiguous()
    dev = g.device
    with torch.cuda.device(dev):
        rc = lib.gsr_densification_stats(P, radii.contiguous().data_ptr() if P else None, _lib.ptr(g), 3,
                                         _lib.ptr(xyz_gradient_accum), _lib.ptr(denom),
                                         _lib.ptr(max_radii2D) if max_radii2D is not None else None,
                                         C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    _lib.check(rc, "gsr_densification_stats")
