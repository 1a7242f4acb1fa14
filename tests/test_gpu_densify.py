"""densify_and_prune on the GPU: scorp_b200.optim.densify_and_prune against
  * tests/golden/densify.npz -- the reference's own GaussianModel.densify_and_prune (gs3dgs and gs2dgs) run on CPU,
  * the numpy oracle on seeded inputs,
  * the reference's sequence of torch calls (gaussian_model.py:460-586) replayed on the same GPU at training size,
    where every decision and every value except the bmm'd child positions must be bit-identical."""
import os
import sys

import numpy as np
import pytest
import torch

import helpers as h
from oracle import optim_oracle as O

sys.path.insert(0, os.path.join(h.ROOT, "tools"))
from torch_densify_sequence import ATTR, model as _model, torch_sequence as _torch_sequence  # noqa: E402

pytestmark = pytest.mark.gpu
G = np.load(os.path.join(os.path.dirname(__file__), "golden", "densify.npz"))


def _cuda(d):
    return {k: torch.from_numpy(np.ascontiguousarray(v)).cuda() for k, v in d.items()}


@pytest.mark.parametrize("c", [str(x) for x in G["cases"]])
def test_golden_reference_model(c):
    from scorp_b200 import optim
    params = _cuda({n: G[f"{c}_in_{n}"] for n in O.GROUPS})
    moments = {n: (torch.from_numpy(G[f"{c}_in_m_{n}"]).cuda(), torch.from_numpy(G[f"{c}_in_v_{n}"]).cuda())
               for n in O.GROUPS if f"{c}_in_m_{n}" in G}
    max_grad, min_opacity, extent, mss, percent_dense = (float(x) for x in G[f"{c}_cfg"])
    out, out_m, counts = optim.densify_and_prune_tensors(
        params, moments, torch.from_numpy(G[f"{c}_in_accum"]).cuda(), torch.from_numpy(G[f"{c}_in_denom"]).cuda(), max_grad,
        min_opacity, extent, mss or None, percent_dense, normals=torch.from_numpy(G[f"{c}_normals"]).cuda())
    assert 2 * counts[3] == G[f"{c}_normals"].shape[0]
    for n in O.GROUPS:
        ref = G[f"{c}_out_{n}"]
        got = out[n].cpu().numpy()
        assert got.shape == ref.shape, (n, got.shape, ref.shape)
        if n in ("xyz", "scaling"):      # children recomputed: CUDA expf/logf vs the CPU run's, bmm order
            np.testing.assert_allclose(got, ref, rtol=2e-6, atol=2e-6)
        else:
            assert np.array_equal(got, ref), n
        if f"{c}_out_m_{n}" in G:
            assert np.array_equal(out_m[n][0].cpu().numpy(), G[f"{c}_out_m_{n}"]), n
            assert np.array_equal(out_m[n][1].cpu().numpy(), G[f"{c}_out_v_{n}"]), n
        else:
            assert n not in out_m


@pytest.mark.parametrize("P,sh_rest,sd,mss", [(300_000, 15, 3, 20), (100_000, 0, 3, None), (150_000, 8, 2, 20)])
def test_matches_torch_sequence_on_gpu(P, sh_rest, sd, mss):
    from scorp_b200 import optim
    pc = _model(P, sh_rest, sd, seed=P % 97)
    par = {n: getattr(pc, a).detach() for n, a in ATTR.items()}
    mom = {n: (pc.optimizer.state[getattr(pc, a)]["exp_avg"], pc.optimizer.state[getattr(pc, a)]["exp_avg_sq"])
           for n, a in ATTR.items() if par[n].numel()}
    args = (0.0002, 0.005, 4.0, mss, 0.01)
    torch.manual_seed(123)
    ref, ref_m = _torch_sequence(par, mom, pc.xyz_gradient_accum, pc.denom, *args)
    torch.manual_seed(123)                # the same generator state gives the same samples, as in the reference
    out, out_m, counts = optim.densify_and_prune_tensors(par, mom, pc.xyz_gradient_accum, pc.denom, *args)
    assert counts[0] > 0 and counts[1] > 0 and counts[2] > 0 and counts[3] >= counts[2]
    assert out["xyz"].shape[0] == counts[0] + counts[1] + 2 * counts[2] == ref["xyz"].shape[0]
    for n in par:
        assert out[n].shape == ref[n].shape, n
        if n == "xyz":                    # children pass through torch.bmm (cuBLAS accumulation order)
            assert torch.allclose(out[n], ref[n], rtol=1e-6, atol=1e-6)
            assert torch.equal(out[n][:counts[0] + counts[1]], ref[n][:counts[0] + counts[1]])
        else:
            assert torch.equal(out[n], ref[n]), (n, int((out[n] != ref[n]).sum()))
        if n in mom:
            assert torch.equal(out_m[n][0], ref_m[n][0]) and torch.equal(out_m[n][1], ref_m[n][1]), n


def test_against_oracle_seeded():
    from scorp_b200 import optim
    pc = _model(40_000, 3, 3, seed=5)
    par = {n: getattr(pc, a).detach() for n, a in ATTR.items()}
    mom = {n: (pc.optimizer.state[getattr(pc, a)]["exp_avg"], pc.optimizer.state[getattr(pc, a)]["exp_avg_sq"]) for n, a in ATTR.items()}
    args = (0.0002, 0.01, 5.0, 20, 0.01)
    out, out_m, counts = optim.densify_and_prune_tensors(par, mom, pc.xyz_gradient_accum, pc.denom, *args,
                                                         generator=torch.Generator(device="cuda").manual_seed(9))
    z = torch.empty((2 * counts[3], 3), device="cuda").normal_(generator=torch.Generator(device="cuda").manual_seed(9))
    ref, ref_m = O.densify_and_prune({n: t.cpu().numpy() for n, t in par.items()},
                                     {n: (m.cpu().numpy(), v.cpu().numpy()) for n, (m, v) in mom.items()},
                                     pc.xyz_gradient_accum.cpu().numpy(), pc.denom.cpu().numpy(), *args, z.cpu().numpy())
    for n in par:
        got = out[n].cpu().numpy()
        # a decision can flip only where numpy's exp differs from CUDA's in the last bit AT a threshold
        assert got.shape == ref[n].shape, (n, got.shape, ref[n].shape)
        if n in ("xyz", "scaling"):
            np.testing.assert_allclose(got, ref[n], rtol=2e-6, atol=2e-6)
        else:
            assert np.array_equal(got, ref[n]), n
        assert np.array_equal(out_m[n][0].cpu().numpy(), ref_m[n][0]) and np.array_equal(out_m[n][1].cpu().numpy(), ref_m[n][1])


def test_model_wrapper_moves_optimizer_state():
    from scorp_b200 import optim
    pc = _model(20_000, 15, 3, seed=3)
    old = {n: getattr(pc, a) for n, a in ATTR.items()}
    steps = {n: pc.optimizer.state[p]["step"] for n, p in old.items()}
    counts = optim.densify_and_prune(pc, 0.0002, 0.005, 4.0, 20)
    rows = counts[0] + counts[1] + 2 * counts[2]
    assert rows != 20_000
    for g in pc.optimizer.param_groups:
        p = g["params"][0]
        n = g["name"]
        assert p is getattr(pc, ATTR[n]) and isinstance(p, torch.nn.Parameter) and p.requires_grad and p.shape[0] == rows
        assert old[n] not in pc.optimizer.state
        st = pc.optimizer.state[p]
        assert st["step"] is steps[n] and st["exp_avg"].shape == p.shape and st["exp_avg_sq"].shape == p.shape
        assert float(st["exp_avg"][counts[0]:].abs().max()) == 0.0          # clones and children start from zero moments
        assert float(st["exp_avg"][:counts[0]].abs().max()) > 0.0
    assert pc.xyz_gradient_accum.shape == (rows, 1) and pc.denom.shape == (rows, 1) and pc.max_radii2D.shape == (rows,)
    assert float(pc.xyz_gradient_accum.abs().max()) == 0 and float(pc.max_radii2D.abs().max()) == 0
    for a in ATTR.values():               # training goes on
        getattr(pc, a).grad = torch.ones_like(getattr(pc, a))
    pc.optimizer.step()


def test_edges():
    from scorp_b200 import optim, _lib
    pc = _model(5_000, 2, 3, seed=8, with_state=False)
    par = {n: getattr(pc, a).detach() for n, a in ATTR.items()}
    # nothing selected, nothing pruned: the identity
    out, out_m, counts = optim.densify_and_prune_tensors(par, {}, pc.xyz_gradient_accum, pc.denom, 1e9, -1.0, 1e9, None, 0.01)
    assert counts == (5_000, 0, 0, 0) and out_m == {}
    assert all(torch.equal(out[n], par[n]) for n in par)
    # everything pruned by opacity
    out, _, counts = optim.densify_and_prune_tensors(par, {}, pc.xyz_gradient_accum, pc.denom, 1e9, 2.0, 1e9, None, 0.01)
    assert counts[:3] == (0, 0, 0) and out["xyz"].shape == (0, 3) and out["f_rest"].shape == (0, 2, 3)
    # empty model
    z = {n: t[:0] for n, t in par.items()}
    out, _, counts = optim.densify_and_prune_tensors(z, {}, pc.denom[:0], pc.denom[:0], 0.0002, 0.005, 4.0, 20, 0.01)
    assert counts == (0, 0, 0, 0) and out["rotation"].shape == (0, 4)
    with pytest.raises(_lib.GsrError):
        optim.densify_and_prune_tensors({n: t.cpu() for n, t in par.items()}, {}, pc.xyz_gradient_accum.cpu(), pc.denom.cpu(),
                                        0.0002, 0.005, 4.0, 20, 0.01)
    with pytest.raises(RuntimeError, match="normals"):
        optim.densify_and_prune_tensors(par, {}, pc.xyz_gradient_accum, pc.denom, 0.0002, 0.005, 4.0, 20, 0.01,
                                        normals=torch.zeros(3, 3, device="cuda"))


def test_reset_opacity():
    """optim.reset_opacity == GaussianModel.reset_opacity: the reference-model golden, the torch expressions on the same
    GPU bit for bit, and the optimizer surgery of replace_tensor_to_optimizer."""
    from scorp_b200 import optim
    pc = _model(50_001, 1, 3, seed=4)
    pc._opacity.data.mul_(1.7)
    old, st_old = pc._opacity, pc.optimizer.state[pc._opacity]
    step = st_old["step"]
    x = torch.sigmoid(old.detach())
    ref = torch.log(torch.min(x, torch.ones_like(x) * 0.01) / (1 - torch.min(x, torch.ones_like(x) * 0.01)))
    new = optim.reset_opacity(pc)
    assert new is pc._opacity and new is not old and isinstance(new, torch.nn.Parameter) and new.requires_grad
    assert torch.equal(new.detach(), ref)
    group = [g for g in pc.optimizer.param_groups if g["name"] == "opacity"][0]
    assert group["params"][0] is new and old not in pc.optimizer.state
    st = pc.optimizer.state[new]
    assert st["step"] is step and float(st["exp_avg"].abs().max()) == 0 and float(st["exp_avg_sq"].abs().max()) == 0
    # golden from the reference's own model (CPU run): same values up to the CPU/CUDA exp, log
    pc2 = _model(900, 1, 3, seed=1, with_state=False)
    pc2._opacity = torch.nn.Parameter(torch.from_numpy(G["reset_in_opacity"]).cuda())
    pc2.optimizer.param_groups[3]["params"][0] = pc2._opacity
    optim.reset_opacity(pc2)
    np.testing.assert_allclose(pc2._opacity.detach().cpu().numpy(), G["reset_out_opacity"], rtol=2e-6, atol=2e-6)
    assert np.array_equal(pc2._opacity.detach().cpu().numpy(), O.reset_opacity(G["reset_in_opacity"])) or \
        np.allclose(pc2._opacity.detach().cpu().numpy(), O.reset_opacity(G["reset_in_opacity"]), rtol=2e-6, atol=2e-6)
