"""Golden vectors for the loss row from the REFERENCE ITSELF: imports
/root/reference/gs3dgs/utils/loss_utils.py (pure torch) in this container, runs l1_loss / ssim and autograd on CPU
fp32 over seeded images and stores inputs, outputs and gradients under tests/golden/l*.npz.
Run here (needs /root/reference); the GPU box only reads the committed .npz files.

    python tools/make_golden_loss.py
"""
import importlib.util
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
REF = Path("/root/reference/gs3dgs/utils/loss_utils.py")

CASES = {            # name: (C, H, W, seed, kind)
    "l1": (3, 37, 53, 11, "noisy"),          # ragged: no multiple of any tile size
    "l2": (3, 8, 150, 12, "noisy"),          # shorter than the window
    "l3": (1, 9, 7, 13, "noisy"),            # smaller than the window both ways
    "l4": (3, 96, 131, 14, "smooth"),        # render-like: smooth image vs slightly different target
    "l5": (3, 40, 40, 15, "identical"),      # ssim = 1, l1 = 0, sign(0) = 0
}


def images(C, H, W, seed, kind):
    g = torch.Generator().manual_seed(seed)
    if kind == "smooth":
        yy, xx = torch.meshgrid(torch.linspace(0, 3, H), torch.linspace(0, 4, W), indexing="ij")
        base = torch.stack([0.5 + 0.4 * torch.sin(xx * (c + 1) + yy) * torch.cos(yy * (c + 2)) for c in range(C)])
        a = (base + 0.02 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
        b = (base + 0.05 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
    else:
        a = torch.rand(C, H, W, generator=g)
        b = (a + 0.3 * torch.randn(C, H, W, generator=g)).clamp(0, 1)
    if kind == "identical":
        b = a.clone()
    return a.float().contiguous(), b.float().contiguous()


def main():
    spec = importlib.util.spec_from_file_location("ref_loss_utils", REF)
    lu = importlib.util.module_from_spec(spec); spec.loader.exec_module(lu)
    torch.set_num_threads(1)
    lam = 0.2
    out = ROOT / "tests" / "golden"
    for name, (C, H, W, seed, kind) in CASES.items():
        a, b = images(C, H, W, seed, kind)
        res = {}
        for key, fn in (("l1", lambda x: lu.l1_loss(x, b)), ("ssim", lambda x: lu.ssim(x, b)),
                        ("loss", lambda x: (1.0 - lam) * lu.l1_loss(x, b) + lam * (1.0 - lu.ssim(x, b)))):
            x = a.clone().requires_grad_(True)
            v = fn(x)
            v.backward()
            res[key] = np.float32(v.item()); res["d" + key] = x.grad.numpy().copy()
        np.savez_compressed(out / f"{name}.npz", img=a.numpy(), gt=b.numpy(), lambda_dssim=np.float32(lam), **res)
        print(name, (C, H, W), kind, {k: float(res[k]) for k in ("l1", "ssim", "loss")})


if __name__ == "__main__":
    sys.exit(main())
