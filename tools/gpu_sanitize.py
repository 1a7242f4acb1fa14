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
torch.cuda.synchronize()
print("done")
