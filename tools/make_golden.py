"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA build (oracle/_ref, compiled
from /root/reference by oracle/Makefile) on the GPU box:

    gpurun -- 'python tools/make_golden.py gpurun_out/golden'   then   cp gpurun_out/golden/*.npz tests/golden/

(g*.npz: 3DGS, s*.npz: 2DGS surfels, k*.npz: simple_knn; `... gpurun_out/golden dsr|knn|dgr` regenerates one family.)
Each file holds the inputs, every intermediate the reference keeps in its scratch buffers, the outputs,
and the gradients as the float64 mean of 8 reference backward runs (its float atomics are order
dependent) together with the per-element spread of those runs.
"""
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import numpy as np
import torch

import helpers as h
from oracle import ref_cuda
from scorp_b200 import synth

CASES = {
    # name: (config, kwargs for make_scene, mode)
    "g1_sh3": ("C2", dict(P=3000, W=160, H=96, seed=11), "sh"),
    "g2_sh0_bg": ("C1", dict(P=2000, W=128, H=128, seed=12, random_bg=True), "sh"),
    "g3_precomp": ("C1", dict(P=1500, W=100, H=70, seed=13, random_bg=True), "precomp"),
    "g4_orbit_sh2": ("C2", dict(P=2500, W=144, H=80, seed=14, sh_degree=2, n_views=3), "sh_view2"),
}


def precomp_inputs(sc):
    """colors_precomp and cov3D_precomp built like the reference's Python paths
    (gs3dgs/gaussian_renderer/__init__.py:88-93, gs3dgs/scene/gaussian_model.py:31-35)."""
    g = torch.Generator().manual_seed(99)
    colors = torch.rand(sc.P, 3, generator=g)
    r, x, y, z = sc.rotations.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    L = R * sc.scales[:, None, :]
    S = L @ L.transpose(1, 2)
    cov = torch.stack([S[:, 0, 0], S[:, 0, 1], S[:, 0, 2], S[:, 1, 1], S[:, 1, 2], S[:, 2, 2]], 1)
    return colors.contiguous(), cov.contiguous()


def main(outdir):
    out = Path(outdir); out.mkdir(parents=True, exist_ok=True)
    for name, (cfg, kw, mode) in CASES.items():
        sc = synth.make_scene(cfg, **kw)
        cam_i = 2 if mode == "sh_view2" else 0
        colors = cov = None
        if mode == "precomp":
            colors, cov = precomp_inputs(sc)
        cpu = dict(means3D=sc.means3D.numpy(), scales=sc.scales.numpy(), rotations=sc.rotations.numpy(),
                   opacities=sc.opacities.numpy(), shs=sc.shs.numpy(), bg=sc.bg.numpy(),
                   viewmatrix=sc.cameras[cam_i].viewmatrix.numpy(), projmatrix=sc.cameras[cam_i].projmatrix.numpy(),
                   campos=sc.cameras[cam_i].campos.numpy())
        if colors is not None:
            cpu["colors_precomp"] = colors.numpy(); cpu["cov3D_precomp"] = cov.numpy()
        sc.to("cuda"); cam = sc.cameras[cam_i]
        cg = None if colors is None else colors.cuda(); vg = None if cov is None else cov.cuda()
        f = h.ref_forward(sc, cam, colors_precomp=cg, cov3D_precomp=vg)
        it = ref_cuda.dgr_intermediates(f)
        grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, kw["seed"])]
        runs = [h.ref_backward(sc, cam, f, grads) for _ in range(8)]
        d = dict(cpu)
        d.update(meta=np.array([sc.P, sc.sh_degree, cam.W, cam.H, f["num_rendered"]], np.int64),
                 tanfov=np.array([cam.tanfovx, cam.tanfovy], np.float64),
                 color=f["color"].cpu().numpy(), depth=f["depth"].cpu().numpy(), alpha=f["alpha"].cpu().numpy(),
                 radii=f["radii"].cpu().numpy())
        for k, v in it.items():
            d["i_" + k] = v.cpu().numpy()
        for k, g in zip(("dL_dcolor", "dL_ddepth", "dL_dalpha"), grads):
            d[k] = g.cpu().numpy()
        for k in runs[0]:
            st = torch.stack([r[k].double() for r in runs])
            d["g_" + k] = st.mean(0).cpu().numpy()
            d["s_" + k] = (st.max(0).values - st.min(0).values).float().cpu().numpy()
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, "P", sc.P, "N", f["num_rendered"], "visible", int((f["radii"] > 0).sum()))


SURFEL_CASES = {
    "s1_sh3": (dict(P=3000, W=160, H=96, seed=21), "sh", 0),
    "s2_sh0_bg": (dict(P=2000, W=128, H=128, seed=22, sh_degree=0, random_bg=True), "sh", 0),
    "s3_precomp": (dict(P=1500, W=100, H=70, seed=23, random_bg=True), "precomp", 0),
    "s4_orbit_sh2": (dict(P=2500, W=144, H=80, seed=24, sh_degree=2, n_views=3), "sh", 2),
}


def surfel_precomp_inputs(sc, cam):
    """colors_precomp and a transMat_precomp built like gs2dgs/gaussian_renderer/__init__.py:78-89
    (splat2world -> world2pix), float32 torch on the CPU."""
    g = torch.Generator().manual_seed(98)
    colors = torch.rand(sc.P, 3, generator=g)
    q = sc.rotations / sc.rotations.norm(dim=1, keepdim=True)
    r, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(-1, 3, 3)
    S = torch.cat([sc.scales, torch.ones_like(sc.scales[:, :1])], 1)
    L = R * S[:, None, :]
    s2w = torch.zeros(sc.P, 4, 4); s2w[:, :3, :3] = L; s2w[:, :3, 3] = sc.means3D; s2w[:, 3, 3] = 1
    s2w = s2w.permute(0, 2, 1)
    W, H = cam.W, cam.H
    ndc2pix = torch.tensor([[W / 2, 0, 0, (W - 1) / 2], [0, H / 2, 0, (H - 1) / 2], [0, 0, 100.0 - 0.01, 0.01],
                            [0, 0, 0, 1]]).float().T
    world2pix = cam.projmatrix.cpu() @ ndc2pix
    tm = (s2w[:, [0, 1, 3]] @ world2pix[:, [0, 1, 3]]).permute(0, 2, 1).reshape(-1, 9)
    return colors.contiguous(), tm.contiguous()


def surfel_main(out):
    for name, (kw, mode, cam_i) in SURFEL_CASES.items():
        sc = synth.make_scene("C4", **kw)
        colors = tm = None
        if mode == "precomp":
            colors, tm = surfel_precomp_inputs(sc, sc.cameras[cam_i])
        cpu = dict(means3D=sc.means3D.numpy(), scales=sc.scales.numpy(), rotations=sc.rotations.numpy(),
                   opacities=sc.opacities.numpy(), shs=sc.shs.numpy(), bg=sc.bg.numpy(),
                   viewmatrix=sc.cameras[cam_i].viewmatrix.numpy(), projmatrix=sc.cameras[cam_i].projmatrix.numpy(),
                   campos=sc.cameras[cam_i].campos.numpy())
        if colors is not None:
            cpu["colors_precomp"] = colors.numpy(); cpu["transMat_precomp"] = tm.numpy()
        sc.to("cuda"); cam = sc.cameras[cam_i]
        cg = None if colors is None else colors.cuda(); tg = None if tm is None else tm.cuda()
        f = h.ref_surfel_forward(sc, cam, colors_precomp=cg, transMat_precomp=tg)
        it = ref_cuda.dsr_intermediates(f)
        grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, kw["seed"], channels=(3, 7))]
        runs = [h.ref_surfel_backward(sc, cam, f, grads) for _ in range(8)]
        d = dict(cpu)
        d.update(meta=np.array([sc.P, sc.sh_degree, cam.W, cam.H, f["num_rendered"]], np.int64),
                 tanfov=np.array([cam.tanfovx, cam.tanfovy], np.float64),
                 color=f["color"].cpu().numpy(), others=f["others"].cpu().numpy(), radii=f["radii"].cpu().numpy())
        vis = (f["radii"] > 0).cpu().numpy()
        for k, v in it.items():
            a = v.cpu().numpy()
            if a.shape[:1] == (sc.P,) and k != "tiles_touched":
                a = a.copy(); a[~vis] = 0          # rows the reference never wrote are uninitialised scratch
            d["i_" + k] = a
        for k, g in zip(("dL_dcolor", "dL_dothers"), grads):
            d[k] = g.cpu().numpy()
        for k in runs[0]:
            st = torch.stack([r[k].double() for r in runs])
            d["g_" + k] = st.mean(0).cpu().numpy()
            d["s_" + k] = (st.max(0).values - st.min(0).values).float().cpu().numpy()
        np.savez_compressed(out / f"{name}.npz", **d)
        print(name, "P", sc.P, "N", f["num_rendered"], "visible", int(vis.sum()))


def knn_main(out):
    g = torch.Generator().manual_seed(31)
    cases = {
        "k1_uniform": torch.rand(6000, 3, generator=g) * 8 - 4,
        "k2_clustered": torch.cat([torch.randn(1500, 3, generator=g) * s + c for s, c in
                                   ((0.05, 0.0), (0.5, 3.0), (2.0, -5.0), (0.001, 1.0))]),
        "k3_planar_dups": torch.cat([torch.cat([torch.rand(2000, 2, generator=g), torch.zeros(2000, 1)], 1),
                                     torch.rand(500, 3, generator=g).repeat(3, 1)]),
        "k4_three_points": torch.rand(3, 3, generator=g),
    }
    for name, pts in cases.items():
        pts = pts.contiguous()
        ref = ref_cuda.knn_dist2(pts.cuda())
        np.savez_compressed(out / f"{name}.npz", points=pts.numpy(), dist2=ref.cpu().numpy())
        print(name, "P", pts.shape[0], "mean", float(ref.mean()))


if __name__ == "__main__":
    outdir = Path(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
    which = sys.argv[2] if len(sys.argv) > 2 else "all"
    outdir.mkdir(parents=True, exist_ok=True)
    if which in ("all", "dgr"):
        main(outdir)
    if which in ("all", "dsr"):
        surfel_main(outdir)
    if which in ("all", "knn"):
        knn_main(outdir)
