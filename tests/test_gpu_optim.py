"""Optimizer row on the GPU: scorp_b200.optim.Adam (one launch, csrc/optim.cu) against
torch.optim.Adam on the same GPU (bit for bit), against tests/golden/adam.npz and the numpy oracle; the fused
densification statistics against the reference's mask-indexing expressions (train_3dgs.py:180-181,
gaussian_model.py:603-605)."""
import numpy as np
import pytest
import torch

import helpers as h

pytestmark = pytest.mark.gpu
NAMES = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
LRS = (0.00016, 0.0025, 0.0025 / 20.0, 0.05, 0.005, 0.001)       # arguments/__init__.py defaults of the reference


def _model(P, seed):
    g = torch.Generator().manual_seed(seed)
    shapes = ((P, 3), (P, 1, 3), (P, 15, 3), (P, 1), (P, 3), (P, 4))
    return [torch.randn(*s, generator=g).cuda() for s in shapes]


def _groups(params):
    return [{"params": [p], "lr": lr, "name": n} for p, lr, n in zip(params, LRS, NAMES)]


def test_adam_bit_identical_to_torch_adam_over_steps():
    from scorp_b200 import optim
    P = 33_333                       # odd sizes: scalar tails, groups that are not multiples of the chunk
    base = _model(P, 1)
    pa = [torch.nn.Parameter(t.clone()) for t in base]
    pb = [torch.nn.Parameter(t.clone()) for t in base]
    oa = torch.optim.Adam(_groups(pa), lr=0.0, eps=1e-15)
    ob = optim.Adam(_groups(pb), lr=0.0, eps=1e-15)
    g = torch.Generator().manual_seed(2)
    for t in range(1, 7):
        if t == 4:                   # update_learning_rate between steps
            oa.param_groups[0]["lr"] = 0.0001; ob.param_groups[0]["lr"] = 0.0001
        for a, b in zip(pa, pb):
            gr = (torch.randn(a.shape, generator=g) * 10.0 ** float(torch.randint(-7, 1, (1,), generator=g))).cuda()
            if t % 2 == 0:
                gr[::3] = 0.0        # invisible Gaussians carry exact zero gradients
            a.grad = gr.clone(); b.grad = gr.clone()
        oa.step(); ob.step()
        for n, a, b in zip(NAMES, pa, pb):
            assert torch.equal(a.data, b.data), (n, t, float((a.data - b.data).abs().max()))
            assert torch.equal(oa.state[a]["exp_avg"], ob.state[b]["exp_avg"]), (n, t)
            assert torch.equal(oa.state[a]["exp_avg_sq"], ob.state[b]["exp_avg_sq"]), (n, t)
    # state dicts interchange (checkpoint written by one, loaded by the other)
    oa2 = torch.optim.Adam(_groups([torch.nn.Parameter(p.data.clone()) for p in pb]), lr=0.0, eps=1e-15)
    oa2.load_state_dict(ob.state_dict())
    assert float(oa2.state[oa2.param_groups[2]["params"][0]]["step"]) == 6.0
    # a parameter without a gradient is skipped, like torch
    pb[0].grad = None
    before = pb[0].data.clone()
    ob.step()
    assert torch.equal(pb[0].data, before)


def test_adam_against_golden_and_oracle():
    from scorp_b200 import optim
    from oracle import optim_oracle as O
    g = np.load(h.ROOT / "tests" / "golden" / "adam.npz")
    params = [torch.nn.Parameter(torch.from_numpy(g[f"p0_{n}"]).cuda()) for n in NAMES]
    opt = optim.Adam([{"params": [p], "lr": float(lr), "name": n} for p, lr, n in zip(params, g["lrs"], NAMES)], lr=0.0, eps=1e-15)
    orc = [(g[f"p0_{n}"], np.zeros_like(g[f"p0_{n}"]), np.zeros_like(g[f"p0_{n}"])) for n in NAMES]
    for t in range(1, int(g["steps"]) + 1):
        if t == 3:
            opt.param_groups[0]["lr"] = 0.00012
        for n, p in zip(NAMES, params):
            p.grad = torch.from_numpy(g[f"g{t}_{n}"]).cuda()
        opt.step()
        for i, (n, p) in enumerate(zip(NAMES, params)):
            lr = 0.00012 if (n == "xyz" and t >= 3) else float(g["lrs"][i])
            orc[i] = O.adam_step(orc[i][0], g[f"g{t}_{n}"], orc[i][1], orc[i][2], lr, t)
            st = opt.state[p]
            for name, ours, ref, oref in (("p", p.data, g[f"p{t}_{n}"], orc[i][0]), ("m", st["exp_avg"], g[f"m{t}_{n}"], orc[i][1]),
                                          ("v", st["exp_avg_sq"], g[f"v{t}_{n}"], orc[i][2])):
                ours = ours.cpu().numpy().astype(np.float64)
                tol = 3e-7 * np.abs(ref) + 3e-7 * float(np.abs(ref).max())      # torch CPU golden: a few ulp (fma or not)
                assert float((np.abs(ours - ref) / tol).max()) <= 1.0, (n, t, name)
                assert np.array_equal(ours.astype(np.float32), oref), (n, t, name, "oracle restates the CUDA sequence exactly")


def test_densification_stats_match_reference_expressions():
    from scorp_b200 import optim
    g = torch.Generator().manual_seed(3)
    P = 100_003
    radii = torch.randint(-3, 60, (P,), generator=g).clamp_min(0).int().cuda()
    vgrad = (torch.randn(P, 3, generator=g) * 1e-4).cuda()
    accum = torch.rand(P, 1, generator=g).cuda(); denom = torch.randint(0, 50, (P, 1), generator=g).float().cuda()
    maxr = torch.randint(0, 40, (P,), generator=g).float().cuda()
    a2, d2, m2 = accum.clone(), denom.clone(), maxr.clone()
    vis = radii > 0                                                 # the reference, verbatim
    maxr[vis] = torch.max(maxr[vis], radii[vis])
    accum[vis] += torch.norm(vgrad[vis, :2], dim=-1, keepdim=True)
    denom[vis] += 1
    optim.add_densification_stats(a2, d2, m2, vgrad, radii)
    assert torch.equal(a2, accum) and torch.equal(d2, denom) and torch.equal(m2, maxr)
    a3 = a2.clone()
    optim.add_densification_stats(a3, d2, None, vgrad, radii)       # max_radii2D optional
    assert torch.equal(a3[~vis], a2[~vis])
    with pytest.raises(RuntimeError):
        optim.add_densification_stats(a2, d2, m2, vgrad[:, :2], radii)


def test_adam_errors():
    from scorp_b200 import optim, _lib
    with pytest.raises(NotImplementedError):
        optim.Adam([torch.nn.Parameter(torch.zeros(3).cuda())], weight_decay=0.1)
    p = torch.nn.Parameter(torch.zeros(3)); p.grad = torch.zeros(3)
    with pytest.raises(_lib.GsrError):
        optim.Adam([p]).step()
