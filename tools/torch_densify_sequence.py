"""The reference's densify_and_prune as its own sequence of torch calls (gs3dgs/scene/gaussian_model.py:460-586;
gs2dgs/scene/gaussian_model.py:447 for the zero third std), replayed on plain CUDA tensors: the same-GPU comparison
target of tests/test_gpu_densify.py and the torch arm of tools/bench_next.py --rows densify. Not part of the product."""
from types import SimpleNamespace

import torch

ATTR = dict(xyz="_xyz", f_dc="_features_dc", f_rest="_features_rest", opacity="_opacity", scaling="_scaling",
            rotation="_rotation")


def model(P, sh_rest, scale_dim, seed, with_state=True, lr=1e-3):
    """a GaussianModel-shaped object with the reference's optimizer layout (gaussian_model.py:192-206)"""
    from scorp_b200 import optim
    g = torch.Generator().manual_seed(seed)
    rn = lambda *s: torch.randn(*s, generator=g)
    pc = SimpleNamespace(percent_dense=0.01)
    pc._xyz = torch.nn.Parameter((rn(P, 3) * 2).cuda())
    pc._features_dc = torch.nn.Parameter(rn(P, 1, 3).cuda())
    pc._features_rest = torch.nn.Parameter((rn(P, sh_rest, 3) * 0.1).cuda())
    pc._opacity = torch.nn.Parameter((rn(P, 1) * 3).cuda())
    pc._scaling = torch.nn.Parameter((rn(P, scale_dim) * 1.2 - 3.0).cuda())
    pc._rotation = torch.nn.Parameter((rn(P, 4) * 1.5).cuda())
    pc.optimizer = optim.Adam([{"params": [getattr(pc, a)], "lr": lr, "name": n} for n, a in ATTR.items()], lr=0.0, eps=1e-15)
    denom = torch.randint(0, 4, (P, 1), generator=g).float()
    pc.denom = denom.cuda()
    pc.xyz_gradient_accum = (torch.rand(P, 1, generator=g) * denom * 4e-4).cuda()
    pc.max_radii2D = torch.randint(0, 60, (P,), generator=g).float().cuda()
    if with_state:
        for a in ATTR.values():
            getattr(pc, a).grad = (rn(*getattr(pc, a).shape) * 1e-3).cuda()
        pc.optimizer.step()
        pc.optimizer.zero_grad(set_to_none=True)
    return pc


def build_rotation(r):          # gs3dgs/utils/general_utils.py:93-115
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3), device=r.device)
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - w * z); R[:, 0, 2] = 2 * (x * z + w * y)
    R[:, 1, 0] = 2 * (x * y + w * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - w * x)
    R[:, 2, 0] = 2 * (x * z - w * y); R[:, 2, 1] = 2 * (y * z + w * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


@torch.no_grad()
def torch_sequence(par, mom, accum, denom, max_grad, min_opacity, extent, max_screen_size, percent_dense, N=2):
    """the reference's rounds with its own torch expressions (gaussian_model.py:460-586), on plain tensors; draws the
    split samples with torch.normal from the current CUDA generator exactly as :511-513 does"""
    par = {k: v.detach().clone() for k, v in par.items()}
    mom = {k: (m.clone(), v.clone()) for k, (m, v) in mom.items()}
    get_scaling = lambda: torch.exp(par["scaling"])

    def cat(new):
        for n in par:
            par[n] = torch.cat((par[n], new[n]), dim=0)
            if n in mom:
                mom[n] = tuple(torch.cat((t, torch.zeros_like(new[n])), dim=0) for t in mom[n])

    def prune(mask):
        for n in par:
            par[n] = par[n][~mask]
            if n in mom:
                mom[n] = tuple(t[~mask] for t in mom[n])

    grads = accum / denom
    grads[grads.isnan()] = 0.0
    sel = torch.where(torch.norm(grads, dim=-1) >= max_grad, True, False)
    sel = torch.logical_and(sel, torch.max(get_scaling(), dim=1).values <= percent_dense * extent)
    cat({n: par[n][sel] for n in par})
    padded = torch.zeros((par["xyz"].shape[0]), device="cuda")
    padded[:grads.shape[0]] = grads.squeeze()
    sel = torch.where(padded >= max_grad, True, False)
    sel = torch.logical_and(sel, torch.max(get_scaling(), dim=1).values > percent_dense * extent)
    stds = get_scaling()[sel].repeat(N, 1)
    if stds.shape[1] == 2:
        stds = torch.cat([stds, 0 * torch.ones_like(stds[:, :1])], dim=-1)
    samples = torch.normal(mean=torch.zeros_like(stds), std=stds)
    rots = build_rotation(par["rotation"][sel]).repeat(N, 1, 1)
    new = {n: par[n][sel].repeat(N, *([1] * (par[n].dim() - 1))) for n in par}
    new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + par["xyz"][sel].repeat(N, 1)
    new["scaling"] = torch.log(get_scaling()[sel].repeat(N, 1) / (0.8 * N))
    cat(new)
    prune(torch.cat((sel, torch.zeros(N * int(sel.sum()), device="cuda", dtype=bool))))
    mask = (torch.sigmoid(par["opacity"]) < min_opacity).squeeze()
    if max_screen_size:
        vs = torch.zeros(par["xyz"].shape[0], device="cuda") > max_screen_size
        ws = get_scaling().max(dim=1).values > 0.1 * extent
        mask = torch.logical_or(torch.logical_or(mask, vs), ws)
    prune(mask)
    return par, mom
