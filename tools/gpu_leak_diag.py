import sys, gc
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from scorp_b200 import synth
import scorp_b200.diff_gaussian_rasterization as dgr
dev = torch.device("cuda", 0)
sc = synth.make_scene("C2", P=200_000).to(dev); cam = sc.cameras[0]
W, H = cam.W, cam.H; hw = H * W
grads = [g.to(dev) for g in synth.upstream_grads(W, H, 1)]
params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
rs0 = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix, cam.projmatrix, sc.sh_degree, cam.campos, False, False)

h_cam = torch.cat([cam.viewmatrix.flatten(), cam.projmatrix.flatten(), cam.campos, sc.bg]).cpu().pin_memory()
h_grads = torch.cat([g.flatten() for g in grads]).cpu().pin_memory()
h_out = torch.empty(1, dtype=torch.float32).pin_memory()

def run(variant):
    for p in params: p.grad = None
    rs = rs0
    if variant in ("slices", "all"):
        d_cam = h_cam.to(dev, non_blocking=True)
        rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d_cam[35:38], 1.0, d_cam[0:16].view(4, 4), d_cam[16:32].view(4, 4), sc.sh_degree, d_cam[32:35], False, False)
    color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(means3D=params[0], means2D=torch.zeros_like(params[0], requires_grad=True), shs=params[1], opacities=params[2], scales=params[3], rotations=params[4])
    if variant == "sum":
        (color.sum() + depth.sum() + alpha.sum()).backward()
    elif variant == "roots":
        torch.autograd.backward([color, depth, alpha], grads)
    elif variant == "color_only":
        torch.autograd.backward([color], [grads[0]])
    elif variant == "fwd_only":
        pass
    elif variant in ("h2d_roots", "all"):
        d_gr = h_grads.to(dev, non_blocking=True)
        gr = [d_gr[:3 * hw].view(3, H, W), d_gr[3 * hw:4 * hw].view(1, H, W), d_gr[4 * hw:].view(1, H, W)]
        torch.autograd.backward([color, depth, alpha], gr)
    elif variant == "slices":
        torch.autograd.backward([color, depth, alpha], grads)
    if variant in ("readback", "all"):
        if variant == "readback":
            torch.autograd.backward([color, depth, alpha], grads)
        h_out.copy_(color.sum().view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

gc.collect(); gc.disable()
for variant in ("roots", "h2d_roots", "slices", "readback", "all"):
    run(variant); run(variant); torch.cuda.synchronize(); base = torch.cuda.memory_allocated()
    for _ in range(5): run(variant)
    torch.cuda.synchronize(); print(variant, "growth per step MB:", (torch.cuda.memory_allocated() - base) / 5 / 1e6)
    gc.collect()
