"""Golden vectors for the optimizer row: torch.optim.Adam configured exactly as
GaussianModel.training_setup does (gs3dgs/scene/gaussian_model.py:192-206: six groups, lr=0.0 default, eps=1e-15) and
the reference's densification-statistics expressions (train_3dgs.py:180-181, gaussian_model.py:603-605), run on CPU
fp32 in this container. Writes tests/golden/adam.npz.   python tools/make_golden_optim.py"""
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
GROUPS = [("xyz", (120, 3), 0.00016), ("f_dc", (120, 1, 3), 0.0025), ("f_rest", (120, 15, 3), 0.0025 / 20.0),
          ("opacity", (120, 1), 0.05), ("scaling", (120, 3), 0.005), ("rotation", (120, 4), 0.001)]
STEPS = 4


def main():
    g = torch.Generator().manual_seed(42)
    params = {n: torch.nn.Parameter(torch.randn(*shape, generator=g)) for n, shape, _ in GROUPS}
    opt = torch.optim.Adam([{"params": [params[n]], "lr": lr, "name": n} for n, _, lr in GROUPS], lr=0.0, eps=1e-15)
    out = {}
    for n in params:
        out[f"p0_{n}"] = params[n].detach().numpy().copy()
    for t in range(1, STEPS + 1):
        if t == 3:      # update_learning_rate (gaussian_model.py:208-214) changes the xyz lr between steps
            opt.param_groups[0]["lr"] = 0.00012
        for n in params:
            scale = 10.0 ** float(torch.randint(-6, 1, (1,), generator=g))
            grad = torch.randn(params[n].shape, generator=g) * scale
            if t == 2:
                grad[::3] = 0.0          # invisible Gaussians: exact zero gradients still move through the moments
            params[n].grad = grad
            out[f"g{t}_{n}"] = grad.numpy().copy()
        opt.step()
        for n in params:
            st = opt.state[params[n]]
            out[f"p{t}_{n}"] = params[n].detach().numpy().copy()
            out[f"m{t}_{n}"] = st["exp_avg"].numpy().copy(); out[f"v{t}_{n}"] = st["exp_avg_sq"].numpy().copy()
    # densification statistics, the reference's expressions verbatim
    P = 2000
    radii = (torch.randint(-2, 40, (P,), generator=g)).clamp_min(0).int()
    vgrad = torch.randn(P, 3, generator=g) * 1e-4
    accum = torch.rand(P, 1, generator=g); denom = torch.randint(0, 50, (P, 1), generator=g).float()
    maxr = torch.randint(0, 30, (P,), generator=g).float()
    out.update(ds_radii=radii.numpy(), ds_vgrad=vgrad.numpy(), ds_accum0=accum.numpy().copy(), ds_denom0=denom.numpy().copy(),
               ds_maxr0=maxr.numpy().copy())
    vis = radii > 0
    maxr[vis] = torch.max(maxr[vis], radii[vis])
    accum[vis] += torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)
    denom[vis] += 1
    out.update(ds_accum1=accum.numpy(), ds_denom1=denom.numpy(), ds_maxr1=maxr.numpy())
    np.savez_compressed(ROOT / "tests" / "golden" / "adam.npz", steps=np.int32(STEPS),
                        lrs=np.array([lr for _, _, lr in GROUPS]), **out)
    print("wrote tests/golden/adam.npz", sum(v.size for v in out.values()), "values")


if __name__ == "__main__":
    main()
