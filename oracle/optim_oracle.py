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
