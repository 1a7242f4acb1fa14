"""Diagnostic: where does the end-to-end (public API + autograd) step spend its time vs the resident C-level step?"""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
from scorp_b200 import synth
import scorp_b200.diff_gaussian_rasterization as dgr
from scorp_b200.diff_gaussian_rasterization import _C

dev = torch.device("cuda", 0)
sc = synth.make_scene("C2").to(dev); cam = sc.cameras[0]
W, H = cam.W, cam.H
grads = [g.to(dev) for g in synth.upstream_grads(W, H, 1)]
params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix, cam.projmatrix,
                                       sc.sh_degree, cam.campos, False, False)
E = torch.Tensor([])

def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(n): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / n * 1e3

def fwd_only():
    with torch.no_grad():
        return dgr.GaussianRasterizer(rs)(params[0], torch.zeros_like(params[0]), params[2], shs=params[1], scales=params[3], rotations=params[4])
def fwd_c():
    return _C.rasterize_gaussians(sc.bg, sc.means3D, E, sc.opacities, sc.scales, sc.rotations, 1.0, E, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, H, W, sc.shs, sc.sh_degree, cam.campos, False, False)
out = fwd_c()
def bwd_c():
    n, color, depth, alpha, radii, geom, binning, img = out
    return _C.rasterize_gaussians_backward(sc.bg, sc.means3D, radii, E, sc.scales, sc.rotations, 1.0, E, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0], grads[1], grads[2], sc.shs, sc.sh_degree, cam.campos, geom, n, binning, img, alpha, False)
def full_autograd(set_none=True):
    if set_none:
        for p in params: p.grad = None
    color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(params[0], torch.zeros_like(params[0], requires_grad=True), params[2], shs=params[1], scales=params[3], rotations=params[4])
    torch.autograd.backward([color, depth, alpha], grads)
def full_autograd_grad():
    color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(params[0], torch.zeros_like(params[0], requires_grad=True), params[2], shs=params[1], scales=params[3], rotations=params[4])
    return torch.autograd.grad([color, depth, alpha], params, grads)
print("fwd C-level        ms", t(fwd_c))
print("fwd API no_grad    ms", t(fwd_only))
print("bwd C-level        ms", t(bwd_c))
print("fwd+bwd autograd   ms", t(full_autograd))
print("fwd+bwd autograd (accumulate into existing .grad) ms", t(lambda: full_autograd(False)))
print("fwd+bwd autograd.grad ms", t(full_autograd_grad))
h = torch.cat([g.flatten() for g in grads]).cpu().pin_memory()
print("H2D 21.8MB pinned  ms", t(lambda: h.to(dev, non_blocking=True)))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
    for _ in range(3): full_autograd()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=18, max_name_column_width=60))
