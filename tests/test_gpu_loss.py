"""Loss row on the GPU: scorp_b200.loss_utils (csrc/loss.cu through the C ABI) against
 (1) tests/golden/l*.npz = the reference's own loss_utils.py outputs and autograd gradients (CPU fp32),
 (2) the CPU oracle at a ragged mid size,
 (3) a torch restatement of loss_utils.py:45-73 running on the same GPU at the full C2 image size,
plus the properties the domain offers (ssim(x, x) = 1, determinism, linearity in the upstream gradient)."""
import glob
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.nn.functional as F

import helpers as h

pytestmark = pytest.mark.gpu
CASES = sorted(glob.glob(str(h.ROOT / "tests" / "golden" / "l?.npz")))

# fp32 tolerance of this row (written here as the task asks): the reference's sigma = E[x^2] - mu^2 cancels in fp32,
# so two fp32 evaluations (its 121-tap conv2d, our 11+11-tap separable pass) agree to ~1e-5 on the scalars and to
# 1e-3 of the largest gradient element; against the double-precision oracle the same bounds hold.
SCALAR_ATOL = 2e-5
GRAD_RTOL_OF_MAX = 1e-3


def grad_close(a, b, n):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.abs(a - b).max()) <= GRAD_RTOL_OF_MAX * float(np.abs(b).max()) + 1e-4 / n


def torch_ssim(img1, img2):
    """loss_utils.py:24-26,39-73 restated in torch (fp32), any device."""
    Cn = img1.size(-3)
    g = torch.tensor([np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    w = g.mm(g.t()).float()[None, None].expand(Cn, 1, 11, 11).contiguous().to(img1.device)
    conv = lambda t: F.conv2d(t, w, padding=5, groups=Cn)
    mu1, mu2 = conv(img1), conv(img2)
    s1 = conv(img1 * img1) - mu1.pow(2); s2 = conv(img2 * img2) - mu2.pow(2); s12 = conv(img1 * img2) - mu1 * mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu1 * mu2 + C1) * (2 * s12 + C2)) / ((mu1.pow(2) + mu2.pow(2) + C1) * (s1 + s2 + C2))
    return m.mean()


@pytest.mark.parametrize("path", CASES, ids=lambda p: Path(p).stem)
def test_against_reference_goldens(path):
    from scorp_b200 import loss_utils as lu
    g = np.load(path)
    lam = float(g["lambda_dssim"])
    b = torch.from_numpy(g["gt"]).cuda()
    n = g["img"].size
    for key, fn in (("l1", lambda x: lu.l1_loss(x, b)), ("ssim", lambda x: lu.ssim(x, b)),
                    ("loss", lambda x: lu.l1_ssim_loss(x, b, lam))):
        x = torch.from_numpy(g["img"]).cuda().requires_grad_(True)
        v = fn(x)
        assert v.dim() == 0
        v.backward()
        assert abs(float(v.detach()) - float(g[key])) <= SCALAR_ATOL, (key, float(v.detach()), float(g[key]))
        assert grad_close(x.grad.cpu().numpy(), g["d" + key], n), \
            (key, float(np.abs(x.grad.cpu().numpy() - g["d" + key]).max()), float(np.abs(g["d" + key]).max()))


def test_against_cpu_oracle_ragged():
    from oracle import gs_oracle as O
    from scorp_b200 import loss_utils as lu
    gen = torch.Generator().manual_seed(77)
    a = torch.rand(3, 101, 67, generator=gen)
    b = (a + 0.15 * torch.randn(3, 101, 67, generator=gen)).clamp(0, 1)
    r = O.l1_ssim(a.numpy(), b.numpy(), 0.35)
    x = a.cuda().requires_grad_(True)
    v = lu.l1_ssim_loss(x, b.cuda(), 0.35)
    v.backward()
    assert abs(float(v.detach()) - r["loss"]) <= SCALAR_ATOL
    assert grad_close(x.grad.cpu().numpy(), r["dloss"], a.numel())
    l1, ss = lu.l1_and_ssim(a.cuda(), b.cuda())
    assert abs(float(l1) - r["l1"]) <= SCALAR_ATOL and abs(float(ss) - r["ssim"]) <= SCALAR_ATOL


def test_full_size_against_torch_on_gpu_and_properties():
    """C2 image size (3 x 840 x 1297): vs the torch restatement on the same GPU; determinism; linearity."""
    from scorp_b200 import loss_utils as lu
    gen = torch.Generator().manual_seed(5)
    H, W = 840, 1297
    yy, xx = torch.meshgrid(torch.linspace(0, 9, H), torch.linspace(0, 14, W), indexing="ij")
    base = torch.stack([0.5 + 0.4 * torch.sin(xx * (c + 1) + yy) * torch.cos(yy * (c + 2)) for c in range(3)])
    a = (base + 0.05 * torch.randn(3, H, W, generator=gen)).clamp(0, 1).cuda()
    b = (base + 0.08 * torch.randn(3, H, W, generator=gen)).clamp(0, 1).cuda()
    lam = 0.2
    x = a.clone().requires_grad_(True)
    ref = (1 - lam) * (x - b).abs().mean() + lam * (1 - torch_ssim(x, b))
    ref.backward()
    y = a.clone().requires_grad_(True)
    out = lu.l1_ssim_loss(y, b, lam)
    (3.0 * out).backward()
    assert abs(float(out) - float(ref)) <= SCALAR_ATOL
    assert grad_close((y.grad / 3.0).cpu().numpy(), x.grad.cpu().numpy(), a.numel())      # linear in the upstream gradient
    # deterministic: no float atomics anywhere
    y2 = a.clone().requires_grad_(True)
    out2 = lu.l1_ssim_loss(y2, b, lam); (3.0 * out2).backward()
    assert torch.equal(out, out2) and torch.equal(y.grad, y2.grad)
    # identities
    assert abs(float(lu.ssim(a, a)) - 1.0) <= 1e-6 and float(lu.l1_loss(a, a)) == 0.0
    z = a.clone().requires_grad_(True)
    lu.l1_ssim_loss(z, a, lam).backward()
    assert float(z.grad.abs().max()) <= 1e-4 / a.numel()


def test_batched_planes_dtype_and_errors():
    from scorp_b200 import loss_utils as lu, _lib
    gen = torch.Generator().manual_seed(9)
    a = torch.rand(2, 3, 33, 47, generator=gen).cuda(); b = torch.rand(2, 3, 33, 47, generator=gen).cuda()
    v4 = lu.ssim(a, b)
    assert abs(float(v4) - float(torch_ssim(a, b))) <= SCALAR_ATOL
    with pytest.raises(_lib.GsrError):
        lu.l1_loss(torch.rand(3, 8, 8), torch.rand(3, 8, 8))
    with pytest.raises(RuntimeError, match="shapes differ"):
        lu.ssim(a, b[:, :, :-1])
    with pytest.raises(NotImplementedError):
        lu.ssim(a, b, window_size=7)
