"""TEST INFRASTRUCTURE (checker): numpy restatement of the optimizer step and the densification statistics that
follow the rasterizer backward in the reference's training loop. Never imported by scorp_b200/.

Follows
  * torch.optim.Adam as the reference configures it -- Adam(l, lr=0.0, eps=1e-15), per-group lr, no weight decay, no
    amsgrad (/root/reference/gs3dgs/scene/gaussian_model.py:192-206; train_3dgs.py:192). The algorithm is torch's
    (torch 2.11 torch/optim/adam.py:_multi_tensor_adam, a third-party dependency of the reference, not vendored in
    /root/reference): m <- lerp(m, g, 1-b1); v <- b2 v + (1-b2) g^2; p <- p - lr/(1-b1^t) * m / (sqrt(v)/sqrt(1-b2^t) + eps).
  * gaussians.max_radii2D[vis] = max(...), add_densification_stats (train_3dgs.py:180-181, gaussian_model.py:603-605).
float32 state, every operation rounded to float32 (fma emulated through float64, exact up to rare double rounding).
Parity status: PINNED by tests/golden/adam.npz = torch.optim.Adam and the reference's mask-indexing expressions run
on CPU fp32 in this container (tools/make_golden_optim.py)."""
from __future__ import annotations

import numpy as np

f32 = np.float32


def _fma(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


def adam_step(p, g, m, v, lr, step, beta1=0.9, beta2=0.999, eps=1e-15):
    """One Adam update of float32 arrays; returns (p, m, v)."""
    p, g, m, v = (np.asarray(t, f32) for t in (p, g, m, v))
    w1, b2, w2 = f32(1.0 - beta1), f32(beta2), f32(1.0 - beta2)
    bc1 = 1.0 - beta1 ** step
    bc2_sqrt = f32((1.0 - beta2 ** step) ** 0.5)
    neg_step = f32((lr / bc1) * -1.0)
    m = _fma(w1, (g - m).astype(f32), m)
    v = _fma(w2, (g * g).astype(f32), (v * b2).astype(f32))
    denom = ((np.sqrt(v).astype(f32) / bc2_sqrt).astype(f32) + f32(eps)).astype(f32)
    p = _fma(neg_step, (m / denom).astype(f32), p)
    return p, m, v


def densification_stats(accum, denom, max_radii, grad2d, radii):
    """Returns updated copies (accum, denom, max_radii) for the Gaussians with radii > 0."""
    accum, denom, max_radii = (np.array(t, f32, copy=True) for t in (accum, denom, max_radii))
    vis = np.asarray(radii) > 0
    gx, gy = np.asarray(grad2d, f32)[:, 0], np.asarray(grad2d, f32)[:, 1]
    nrm = np.sqrt(((gx * gx).astype(f32) + (gy * gy).astype(f32)).astype(f32)).astype(f32)
    accum[vis] = (accum[vis] + nrm[vis]).astype(f32)
    denom[vis] = denom[vis] + f32(1.0)
    max_radii[vis] = np.maximum(max_radii[vis], np.asarray(radii)[vis].astype(f32))
    return accum, denom, max_radii


# ---------------------------------------------------------------------------------------------------------------
# densify_and_prune: the reference's sequence, round by round (no fusion here -- this is the checker).
# Pinned by tests/golden/densify.npz = the reference's own GaussianModel.densify_and_prune run on CPU
# (tools/make_golden_densify.py).
GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def _build_rotation(r):
    """/root/reference/gs3dgs/utils/general_utils.py:93-115 (every torch op rounds to float32)."""
    r = np.asarray(r, f32)
    sq = lambda a: (a * a).astype(f32)
    norm = np.sqrt((((sq(r[:, 0]) + sq(r[:, 1])).astype(f32) + sq(r[:, 2])).astype(f32) + sq(r[:, 3])).astype(f32)).astype(f32)
    q = (r / norm[:, None]).astype(f32)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    mul = lambda a, b: (a * b).astype(f32)
    R = np.zeros((r.shape[0], 3, 3), f32)
    R[:, 0, 0] = f32(1) - f32(2) * (mul(y, y) + mul(z, z)).astype(f32)
    R[:, 0, 1] = f32(2) * (mul(x, y) - mul(w, z)).astype(f32)
    R[:, 0, 2] = f32(2) * (mul(x, z) + mul(w, y)).astype(f32)
    R[:, 1, 0] = f32(2) * (mul(x, y) + mul(w, z)).astype(f32)
    R[:, 1, 1] = f32(1) - f32(2) * (mul(x, x) + mul(z, z)).astype(f32)
    R[:, 1, 2] = f32(2) * (mul(y, z) - mul(w, x)).astype(f32)
    R[:, 2, 0] = f32(2) * (mul(x, z) - mul(w, y)).astype(f32)
    R[:, 2, 1] = f32(2) * (mul(y, z) + mul(w, x)).astype(f32)
    R[:, 2, 2] = f32(1) - f32(2) * (mul(x, x) + mul(y, y)).astype(f32)
    return R


def densify_and_prune(params, moments, accum, denom, max_grad, min_opacity, extent, max_screen_size, percent_dense,
                      normals, N=2):
    """params {name: [P, ...]} stored tensors, moments {name: (exp_avg, exp_avg_sq)} or {} -> (params, moments) after
    GaussianModel.densify_and_prune (/root/reference/gs3dgs/scene/gaussian_model.py:572-586; the 2DGS copy at
    gs2dgs/scene/gaussian_model.py:478-492 appends a zero third std, :447). normals [N * selected, 3] are the standard
    normal samples torch.normal scales by std (:511-513)."""
    par = {n: np.array(params[n], f32, copy=True) for n in GROUPS}
    mom = {n: tuple(np.array(t, f32, copy=True) for t in moments[n]) for n in GROUPS if moments and n in moments}
    sigmoid = lambda x: (f32(1) / (f32(1) + np.exp(-x).astype(f32)).astype(f32)).astype(f32)
    scaling = lambda: np.exp(par["scaling"]).astype(f32)                                    # get_scaling
    thr, dense, big = f32(max_grad), f32(percent_dense * extent), f32(0.1 * extent)

    def cat(new):            # densification_postfix -> cat_tensors_to_optimizer (:460-500)
        for n in GROUPS:
            par[n] = np.concatenate([par[n], new[n].astype(f32)], 0)
            if n in mom:
                mom[n] = tuple(np.concatenate([t, np.zeros_like(new[n], f32)], 0) for t in mom[n])

    def prune(mask):         # prune_points -> _prune_optimizer (:425-458)
        keep = ~mask
        for n in GROUPS:
            par[n] = par[n][keep]
            if n in mom:
                mom[n] = tuple(t[keep] for t in mom[n])

    with np.errstate(divide="ignore", invalid="ignore"):
        grads = (np.asarray(accum, f32).reshape(-1) / np.asarray(denom, f32).reshape(-1)).astype(f32)   # :573
    grads[np.isnan(grads)] = 0.0                                                                          # :574
    # densify_and_clone (:559-570)
    sel = (np.sqrt((grads * grads).astype(f32)).astype(f32) >= thr) & (scaling().max(1) <= dense)
    cat({n: par[n][sel] for n in GROUPS})
    # densify_and_split (:502-527)
    padded = np.zeros(par["xyz"].shape[0], f32); padded[:grads.shape[0]] = grads
    sel = (padded >= thr) & (scaling().max(1) > dense)
    stds = np.tile(scaling()[sel], (N, 1))
    if stds.shape[1] == 2:                                                   # gs2dgs :447
        stds = np.concatenate([stds, np.zeros_like(stds[:, :1])], 1)
    normals = np.asarray(normals, f32)
    assert normals.shape == stds.shape, (normals.shape, stds.shape)
    samples = ((normals * stds).astype(f32) + f32(0)).astype(f32)
    rots = np.tile(_build_rotation(par["rotation"][sel]), (N, 1, 1))
    moved = np.einsum("nij,nj->ni", rots.astype(np.float64), samples.astype(np.float64)).astype(f32)   # torch.bmm
    new = {n: np.tile(par[n][sel], (N,) + (1,) * (par[n].ndim - 1)) for n in GROUPS}
    new["xyz"] = (moved + new["xyz"]).astype(f32)
    inv = f32(1.0) / f32(0.8 * N)                    # tensor / python float = multiply by the float reciprocal
    new["scaling"] = np.log((np.tile(scaling()[sel], (N, 1)) * inv).astype(f32)).astype(f32)
    cat(new)
    prune(np.concatenate([sel, np.zeros(N * int(sel.sum()), bool)]))
    # final prune (:579-584); max_radii2D was zeroed by densification_postfix (:498)
    mask = sigmoid(par["opacity"]).reshape(-1) < f32(min_opacity)
    if max_screen_size:
        mask = mask | (np.zeros(mask.shape, f32) > f32(max_screen_size)) | (scaling().max(1) > big)
    prune(mask)
    return par, mom


def reset_opacity(opacity, cap=0.01):
    """GaussianModel.reset_opacity (/root/reference/gs3dgs/scene/gaussian_model.py:253-256) with inverse_sigmoid
    (gs3dgs/utils/general_utils.py:20-21): log(x / (1 - x)) of min(sigmoid(opacity), 0.01); the moments become zeros."""
    o = np.asarray(opacity, f32)
    a = (f32(1) / (f32(1) + np.exp(-o).astype(f32)).astype(f32)).astype(f32)
    x = np.minimum(a, f32(cap))
    return np.log((x / (f32(1) - x).astype(f32)).astype(f32)).astype(f32)
