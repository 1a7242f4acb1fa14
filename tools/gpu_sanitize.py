"""Small end-to-end pass over every kernel for compute-sanitizer (memcheck / racecheck): 3DGS and 2DGS forward +
backward (plain and factored), mark_visible, distCUDA2, with ragged image sizes and a crowded tile."""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), "..", "tests")); sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
import torch
import helpers as h
from scorp_b200 import synth, multiview as mv
from scorp_b200.simple_knn._C import distCUDA2
from scorp_b200.diff_gaussian_rasterization import _C as c3

for cfg, kw, surfel in (("C1", dict(P=6000, W=200, H=136), False), ("C1", dict(P=30000, W=96, H=64), False),
                        ("C4", dict(P=6000, W=200, H=136), True), ("C2", dict(P=4000, W=130, H=70, sh_degree=1), False)):
    sc = h.make(cfg, **kw); cam = sc.cameras[0]
    fwd_f, bwd_f = (h.ours_surfel_forward, h.ours_surfel_backward) if surfel else (h.ours_forward, h.ours_backward)
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5, channels=(3, 7) if surfel else (3, 1, 1))]
    fo = fwd_f(sc, cam)
    g1 = bwd_f(sc, cam, fo, grads)
    mv.install_factored_exchange(enable=True)
    g2 = bwd_f(sc, cam, fo, grads)
    mv.install_factored_exchange(enable=False)
    torch.cuda.synchronize()
    print(cfg, kw, "N =", fo["num_rendered"], "ok", bool(torch.isfinite(g1["sh"]).all()), bool(torch.isfinite(g2["sh"]).all()))
sc = h.make("C1", P=5000, W=64, H=64)
print("visible", int(c3.mark_visible(sc.means3D, sc.cameras[0].viewmatrix, sc.cameras[0].projmatrix).sum()))
print("knn", float(distCUDA2(torch.rand(10000, 3, device="cuda")).mean()))

# rows next to the boundary: model-space 3DGS / 2DGS, batched views, loss, Adam, densification statistics
import torch.nn.functional as F
from scorp_b200 import model_space as ms, loss_utils as lu, optim
import scorp_b200.diff_gaussian_rasterization as dgr
import scorp_b200.diff_surfel_rasterization as dsr
for cfg, kw, mod in (("C2", dict(P=5000, W=150, H=90), dgr), ("C1", dict(P=3000, W=70, H=50), dgr), ("C4", dict(P=5000, W=150, H=90), dsr)):
    sc = h.make(cfg, **kw); cam = sc.cameras[0]
    rs = mod.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix, cam.projmatrix,
                                           sc.sh_degree, cam.campos, False, False)
    op = sc.opacities.clamp(1e-4, 1 - 1e-4)
    st = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs[:, :1].contiguous(), sc.shs[:, 1:].contiguous(),
                                                   torch.log(op / (1 - op)), torch.log(sc.scales), sc.rotations * 2.0)]
    m2d = torch.zeros_like(st[0], requires_grad=True)
    fn = ms.rasterize_surfel_model if mod is dsr else ms.rasterize_model
    outs = fn(st[0], m2d, st[1], st[2], st[3], st[4], st[5], rs)
    gt = torch.rand_like(outs[0])
    loss = lu.l1_ssim_loss(outs[0], gt, 0.2) + 0.01 * outs[2].sum()
    loss.backward()
    accum = torch.zeros(sc.P, 1, device="cuda"); den = torch.zeros(sc.P, 1, device="cuda"); mr = torch.zeros(sc.P, device="cuda")
    optim.add_densification_stats(accum, den, mr, m2d.grad, outs[1])
    params = [torch.nn.Parameter(t.detach().clone()) for t in st if t.numel()]
    for p_, t in zip(params, [t for t in st if t.numel()]):
        p_.grad = t.grad
    optim.Adam([{"params": [p_], "lr": 1e-3} for p_ in params], lr=0.0, eps=1e-15).step()
    # densify_and_prune at an odd size (the statistics just accumulated decide who is cloned / split)
    names = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
    dn, _, cnt = optim.densify_and_prune_tensors({n_: t.detach() for n_, t in zip(names, st)}, {}, accum * 50, den, 0.0002, 0.005,
                                                 4.0, 20, 0.01)
    torch.cuda.synchronize()
    print("model-space", cfg, kw, "loss", float(loss.detach()), "stats", float(den.sum()), "densify", cnt, tuple(dn["xyz"].shape))
sc = synth.make_scene("C5", P=4000, W=100, H=80, n_views=5, rotations_npz=synth.random_rotations(5)).to("cuda")
cams = sc.cameras
E = torch.Tensor([])
n, col, dep, alp, rad = c3.rasterize_gaussians_views(sc.bg, sc.means3D, E, sc.opacities, sc.scales, sc.rotations, 1.0, E,
                                                     torch.stack([c.viewmatrix for c in cams]), torch.stack([c.projmatrix for c in cams]),
                                                     cams[0].tanfovx, cams[0].tanfovy, 80, 100, sc.shs, sc.sh_degree,
                                                     torch.stack([c.campos for c in cams]), False, False)
torch.cuda.synchronize()
print("views", n, float(col.sum()))
import tempfile
from scorp_b200 import ply_io
with tempfile.TemporaryDirectory() as tmp:
    mdl = dict(xyz=st[0].detach(), features_dc=st[1].detach(), features_rest=st[2].detach(), opacity=st[3].detach(),
               scaling=st[4].detach(), rotation=st[5].detach())
    ply_io.save_ply(mdl, os.path.join(tmp, "m.ply"))
    bk = ply_io.load_ply(os.path.join(tmp, "m.ply"))
    print("ply", tuple(bk["features_rest"].shape), all(torch.equal(bk[k], mdl[k]) for k in mdl))
print("done")
