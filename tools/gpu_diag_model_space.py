"""GPU diagnostic: which activation of the model-space preprocess differs in bits from torch's kernels.
Compares the per-Gaussian records (opacity, cov3D, conic, rgb) of the stock path fed with torch activations against
the model-space path, and torch's activations against candidate float sequences evaluated with torch ops."""
import sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch
import torch.nn.functional as F
import helpers as h
import scorp_b200.diff_gaussian_rasterization as dgr
from scorp_b200 import model_space as ms

sc = h.make("C2", P=200_000, W=640, H=400); cam = sc.cameras[0]
g = torch.Generator().manual_seed(5)
op = sc.opacities.clamp(1e-4, 1 - 1e-4)
raw_op = torch.log(op / (1 - op)); raw_sc = torch.log(sc.scales)
raw_rot = sc.rotations * (0.25 + 3.0 * torch.rand(sc.P, 1, generator=g)).cuda()
rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix, cam.projmatrix,
                                       sc.sh_degree, cam.campos, False, False)
a_op, a_sc, a_rot = torch.sigmoid(raw_op), torch.exp(raw_sc), F.normalize(raw_rot)
sc2 = h.make("C2", P=200_000, W=640, H=400)
sc2.opacities, sc2.scales, sc2.rotations = a_op, a_sc, a_rot
fo = h.ours_forward(sc2, cam)
io = h.ours_intermediates(fo)
out = ms.model_forward(sc.means3D, sc.shs[:, :1].contiguous(), sc.shs[:, 1:].contiguous(), raw_op, raw_sc, raw_rot, rs)
fm = dict(zip(("num_rendered", "color", "depth", "alpha", "radii", "geom", "binning", "image"), out[:8]))
fm.update(P=sc.P, H=cam.H, W=cam.W)
im = h.ours_intermediates(fm)
vis = fo["radii"] > 0
print("visible", int(vis.sum()), "radii equal", torch.equal(fo["radii"], fm["radii"]))
for k in ("means2D", "conic_opacity", "depths", "rgb", "cov3D"):
    a, b = io[k][vis], im[k][vis]
    ne = (a.view(torch.int32) != b.view(torch.int32))
    print(k, "mismatching elements", int(ne.sum()), "of", ne.numel(), "per column", ne.reshape(ne.shape[0], -1).sum(0).tolist())
# candidates, evaluated with torch elementwise kernels (each op rounds once)
x = raw_op.flatten()
c1 = 1.0 / (1.0 + torch.exp(-x))
print("sigmoid == 1/(1+exp(-x)):", int((torch.sigmoid(x).view(torch.int32) != c1.view(torch.int32)).sum()), "differ of", x.numel())
q = raw_rot
sq = q * q
n_pair = torch.sqrt((sq[:, 0] + sq[:, 1]) + (sq[:, 2] + sq[:, 3])).clamp_min(1e-12)
n_seq = torch.sqrt(((sq[:, 0] + sq[:, 1]) + sq[:, 2]) + sq[:, 3]).clamp_min(1e-12)
n_fma = torch.sqrt(torch.addcmul(torch.addcmul(torch.addcmul(sq[:, 0], q[:, 1], q[:, 1]), q[:, 2], q[:, 2]), q[:, 3], q[:, 3])).clamp_min(1e-12)
n_t = q.norm(2, 1)
for name, n in (("pairwise", n_pair), ("sequential", n_seq), ("fma chain", n_fma)):
    print("norm ==", name, ":", int((n.view(torch.int32) != n_t.view(torch.int32)).sum()), "differ")
    d = q / n[:, None]
    print("   normalize == q /", name, ":", int((d.view(torch.int32) != a_rot.view(torch.int32)).sum()), "elements differ")
