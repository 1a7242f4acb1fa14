"""Diagnostic: per-step wall time of bench.py's e2e step variants."""
import sys, time
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
import torch
from scorp_b200 import synth
import scorp_b200.diff_gaussian_rasterization as dgr
dev = torch.device("cuda", 0)
sc = synth.make_scene("C2").to(dev); cam = sc.cameras[0]
W, H = cam.W, cam.H; hw = H * W
grads = [g.to(dev) for g in synth.upstream_grads(W, H, 1)]
params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
h_cam = torch.cat([cam.viewmatrix.flatten(), cam.projmatrix.flatten(), cam.campos, sc.bg]).cpu().pin_memory()
h_grads = torch.cat([g.flatten() for g in grads]).cpu().pin_memory()
h_out = torch.empty(1, dtype=torch.float32).pin_memory()

def step(h2d=True, slices=True, readback=True):
    if h2d:
        d_cam = h_cam.to(dev, non_blocking=True); d_gr = h_grads.to(dev, non_blocking=True)
        gr = [d_gr[:3 * hw].view(3, H, W), d_gr[3 * hw:4 * hw].view(1, H, W), d_gr[4 * hw:].view(1, H, W)]
    else:
        gr = grads
    if h2d and slices:
        rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, d_cam[35:38], 1.0, d_cam[0:16].view(4, 4), d_cam[16:32].view(4, 4), sc.sh_degree, d_cam[32:35], False, False)
    else:
        rs = dgr.GaussianRasterizationSettings(H, W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix, cam.projmatrix, sc.sh_degree, cam.campos, False, False)
    for p in params: p.grad = None
    color, radii, depth, alpha = dgr.GaussianRasterizer(rs)(means3D=params[0], means2D=torch.zeros_like(params[0], requires_grad=True), shs=params[1], opacities=params[2], scales=params[3], rotations=params[4])
    torch.autograd.backward([color, depth, alpha], gr)
    if readback:
        h_out.copy_(color.detach().sum().view(1), non_blocking=True)
        torch.cuda.current_stream().synchronize()

for name, kw in [("full", {}), ("no readback/sync", dict(readback=False)), ("no h2d", dict(h2d=False)), ("h2d no slices", dict(slices=False)), ("full again", {})]:
    ts = []
    for i in range(25):
        torch.cuda.synchronize(); t0 = time.perf_counter(); step(**kw); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
    print(f"{name:20s}", " ".join(f"{x:.2f}" for x in ts))

import gc, statistics
for label, off in (("gc on", False), ("gc off", True), ("gc on again", False)):
    if off:
        gc.collect(); gc.disable()
    ts = []
    for i in range(400):
        torch.cuda.synchronize(); t0 = time.perf_counter(); step(); torch.cuda.synchronize(); ts.append((time.perf_counter() - t0) * 1e3)
        if torch.cuda.memory_allocated() > 20e9:
            print("memory keeps growing:", torch.cuda.memory_allocated() / 1e9, "GB after", i, "steps"); break
    if off:
        gc.enable()
    med = statistics.median(ts)
    spikes = [(i, round(x, 2)) for i, x in enumerate(ts) if x > 1.5 * med]
    print(f"allocated {torch.cuda.memory_allocated() / 1e9:.2f} GB")
    print(f"{label:12s} median {med:.3f} mean {sum(ts)/len(ts):.3f} max {max(ts):.2f} spikes {len(spikes)}: {spikes[:12]}")
