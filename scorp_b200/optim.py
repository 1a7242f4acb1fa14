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

`densify_and_prune(gaussians, max_grad, min_opacity, extent, max_screen_size)` is GaussianModel.densify_and_prune
(gaussian_model.py:572-586; gs2dgs/scene/gaussian_model.py:478-492) with the model as first argument: clone, split,
prune of the six parameter tensors and their Adam moments in one plan + one gather (csrc/densify.cu) instead of four
rounds of mask indexing and torch.cat. `densify_and_prune_tensors` is the same on plain tensors.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib

__all__ = ["Adam", "add_densification_stats", "densify_and_prune", "densify_and_prune_tensors", "reset_opacity"]


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


# optimizer group name -> (GaussianModel attribute, gather kind); gaussian_model.py:193-200
_GROUPS = (("xyz", "_xyz", 1), ("f_dc", "_features_dc", 0), ("f_rest", "_features_rest", 0),
           ("opacity", "_opacity", 0), ("scaling", "_scaling", 2), ("rotation", "_rotation", 0))


def densify_and_prune_tensors(params, moments, xyz_gradient_accum, denom, max_grad, min_opacity, extent,
                              max_screen_size, percent_dense, N=2, normals=None, generator=None):
    """The whole clone / split / prune sequence on plain tensors.

    params:  {"xyz": [P,3], "f_dc": [P,1,3], "f_rest": [P,R,3], "opacity": [P,1], "scaling": [P,3|2], "rotation": [P,4]}
             stored (pre-activation) CUDA fp32 tensors.
    moments: {name: (exp_avg, exp_avg_sq)} for the groups that have optimizer state (others: absent or None).
    normals: [N * n_selected, 3] standard normal samples; drawn with `generator` when None, with the shape the
             reference's torch.normal call draws (gaussian_model.py:511-513), so the same seed gives the same children.
    Returns (new_params, new_moments, counts) with counts = (kept originals, clones, split parents kept, selected)."""
    lib = _lib.lib()
    xyz = params["xyz"]
    if not xyz.is_cuda:
        raise _lib.GsrError("densify_and_prune needs CUDA tensors: there is no CPU path")
    dev, P = xyz.device, xyz.shape[0]
    src = {}
    for name, _, _ in _GROUPS:
        t = params[name]
        if t.dtype != torch.float32 or t.shape[0] != P or t.device != dev:
            raise RuntimeError(f"{name}: expected an fp32 tensor with {P} rows on {dev}")
        src[name] = t.detach().contiguous()
    scale_dim = src["scaling"].shape[1]
    if src["rotation"].shape[1:] != (4,) or src["xyz"].shape[1:] != (3,) or src["opacity"].numel() != P:
        raise RuntimeError("expected xyz [P,3], opacity [P,1], rotation [P,4]")
    acc, den = xyz_gradient_accum.detach().contiguous(), denom.detach().contiguous()
    if acc.numel() != P or den.numel() != P or acc.dtype != torch.float32 or den.dtype != torch.float32:
        raise RuntimeError("xyz_gradient_accum and denom must be fp32 with one element per Gaussian")
    cfg = _lib.GsrDensifyCfg(float(max_grad), float(percent_dense * extent), float(min_opacity), float(0.1 * extent),
                             float(max_screen_size) if max_screen_size else 0.0, 1 if max_screen_size else 0, int(N))
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    counts = (C.c_int32 * 4)()
    with torch.cuda.device(dev):
        ws = torch.empty(max(int(lib.gsr_densify_workspace_bytes(P, int(N))), 16), dtype=torch.uint8, device=dev)
        _lib.check(lib.gsr_densify_plan(P, scale_dim, _lib.ptr(src["scaling"]), _lib.ptr(src["opacity"]), _lib.ptr(acc),
                                        _lib.ptr(den), C.byref(cfg), _lib.ptr(ws), counts, stream), "gsr_densify_plan")
        n_keep, n_clone, n_child, n_sel = (int(c) for c in counts)
        rows = n_keep + n_clone + int(N) * n_child
        if normals is None:
            normals = torch.empty((int(N) * n_sel, 3), device=dev).normal_(generator=generator)
        if normals.shape != (int(N) * n_sel, 3) or normals.dtype != torch.float32 or normals.device != dev:
            raise RuntimeError(f"normals: expected fp32 [{int(N) * n_sel}, 3] on {dev}")
        normals = normals.contiguous()
        arr = (_lib.GsrDensifyGroup * len(_GROUPS))()
        out, out_m, keep = {}, {}, []
        for a, (name, _, kind) in zip(arr, _GROUPS):
            t = src[name]
            out[name] = torch.empty((rows,) + tuple(t.shape[1:]), device=dev)
            mv = (moments or {}).get(name)
            a.param, a.out_param = _lib.ptr(t), _lib.ptr(out[name])
            a.width, a.kind = (t.numel() // P if P else 0), kind
            if mv is not None:
                m, v = (x.detach().contiguous() for x in mv)
                if m.shape != t.shape or v.shape != t.shape:
                    raise RuntimeError(f"{name}: optimizer moments do not have the parameter's shape")
                out_m[name] = (torch.empty_like(out[name]), torch.empty_like(out[name]))
                a.exp_avg, a.exp_avg_sq = _lib.ptr(m), _lib.ptr(v)
                a.out_exp_avg, a.out_exp_avg_sq = _lib.ptr(out_m[name][0]), _lib.ptr(out_m[name][1])
                keep.append((m, v))          # the contiguous copies live until the launch is queued
        live = [g for g in arr if g.width > 0 and rows > 0]
        if P and rows:
            packed = (_lib.GsrDensifyGroup * len(live))(*live)
            _lib.check(lib.gsr_densify_apply(P, scale_dim, C.byref(cfg), _lib.ptr(ws), counts, len(live), packed,
                                             _lib.ptr(src["rotation"]), _lib.ptr(src["scaling"]), _lib.ptr(normals),
                                             stream), "gsr_densify_apply")
    return out, out_m, (n_keep, n_clone, n_child, n_sel)


@torch.no_grad()
def densify_and_prune(gaussians, max_grad, min_opacity, extent, max_screen_size, N=2, generator=None, empty_cache=True):
    """GaussianModel.densify_and_prune (gaussian_model.py:572-586) on a GaussianModel-shaped object: replaces the six
    nn.Parameters and the optimizer's per-group state exactly as cat_tensors_to_optimizer / _prune_optimizer leave them
    (new nn.Parameter per group, state moved to it, `step` kept), and zeroes xyz_gradient_accum, denom, max_radii2D
    at the new size (densification_postfix :496-498)."""
    opt = gaussians.optimizer
    by_name = {g["name"]: g for g in opt.param_groups}
    params, moments = {}, {}
    for name, attr, _ in _GROUPS:
        params[name] = getattr(gaussians, attr)
        st = opt.state.get(by_name[name]["params"][0], None) if name in by_name else None
        if st is not None and "exp_avg" in st:
            moments[name] = (st["exp_avg"], st["exp_avg_sq"])
    out, out_m, counts = densify_and_prune_tensors(params, moments, gaussians.xyz_gradient_accum, gaussians.denom, max_grad,
                                                   min_opacity, extent, max_screen_size, gaussians.percent_dense, N=N,
                                                   generator=generator)
    for name, attr, _ in _GROUPS:
        new = torch.nn.Parameter(out[name].requires_grad_(True))
        if name in by_name:
            group = by_name[name]
            old = group["params"][0]
            st = opt.state.get(old, None)
            if st is not None:
                if name in out_m:
                    st["exp_avg"], st["exp_avg_sq"] = out_m[name]
                del opt.state[old]
                opt.state[new] = st
            group["params"][0] = new
        setattr(gaussians, attr, new)
    rows, dev = out["xyz"].shape[0], out["xyz"].device
    gaussians.xyz_gradient_accum = torch.zeros((rows, 1), device=dev)
    gaussians.denom = torch.zeros((rows, 1), device=dev)
    gaussians.max_radii2D = torch.zeros((rows,), device=dev)
    if empty_cache:
        torch.cuda.empty_cache()          # as the reference does (:586)
    return counts


@torch.no_grad()
def reset_opacity(gaussians, cap: float = 0.01):
    """GaussianModel.reset_opacity (gaussian_model.py:253-256): opacity <- inverse_sigmoid(min(sigmoid(opacity), 0.01)),
    and, as replace_tensor_to_optimizer does (:410-423), a new nn.Parameter in the "opacity" group with zeroed moments
    (`step` kept). One launch; the values are written into a fresh tensor's storage like the reference's new tensor."""
    lib = _lib.lib()
    old = gaussians._opacity
    if not old.is_cuda:
        raise _lib.GsrError("reset_opacity needs CUDA tensors: there is no CPU path")
    new = old.detach().clone(memory_format=torch.contiguous_format)
    group = next((g for g in gaussians.optimizer.param_groups if g.get("name") == "opacity"), None)
    st = gaussians.optimizer.state.get(group["params"][0], None) if group is not None else None
    m = v = None
    if st is not None and "exp_avg" in st:
        m, v = torch.empty_like(new), torch.empty_like(new)
    dev = new.device
    with torch.cuda.device(dev):
        rc = lib.gsr_reset_opacity(new.numel(), _lib.ptr(new), _lib.ptr(m) if m is not None else None,
                                   _lib.ptr(v) if v is not None else None, float(cap),
                                   C.c_void_p(torch.cuda.current_stream(dev).cuda_stream))
    _lib.check(rc, "gsr_reset_opacity")
    param = torch.nn.Parameter(new.requires_grad_(True))
    if group is not None:
        if st is not None:
            if m is not None:
                st["exp_avg"], st["exp_avg_sq"] = m, v
            del gaussians.optimizer.state[group["params"][0]]
            gaussians.optimizer.state[param] = st
        group["params"][0] = param
    gaussians._opacity = param
    return param
