"""Generates tests/golden/*.npz by running the UNMODIFIED reference CUDA build (oracle/_ref, compiled
from /root/reference by oracle/Makefile) on the GPU box:

    gpurun -- 'python tools/make_golden.py gpurun_out/golden'   then   cp gpurun_out/golden/*.npz tests/golden/

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


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else "gpurun_out/golden")
