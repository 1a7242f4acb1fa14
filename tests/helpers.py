"""Shared test helpers: run the product (through its reference-shaped `_C` level, i.e. the C ABI),
the reference CUDA build (oracle/_ref via oracle/ref_cuda.py) and compare."""
from __future__ import annotations

import ctypes as C
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

from scorp_b200 import _lib, synth  # noqa: E402


def empty():
    return torch.Tensor([])


def ours_forward(scene, cam, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0, debug=False):
    from scorp_b200.diff_gaussian_rasterization import _C
    shs = empty() if colors_precomp is not None else scene.shs
    col = colors_precomp if colors_precomp is not None else empty()
    sc = empty() if cov3D_precomp is not None else scene.scales
    rot = empty() if cov3D_precomp is not None else scene.rotations
    cov = cov3D_precomp if cov3D_precomp is not None else empty()
    out = _C.rasterize_gaussians(scene.bg, scene.means3D, col, scene.opacities, sc, rot, scale_modifier, cov,
                                 cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, shs,
                                 scene.sh_degree, cam.campos, False, debug)
    keys = ("num_rendered", "color", "depth", "alpha", "radii", "geom", "binning", "image")
    d = dict(zip(keys, out))
    d.update(P=scene.P, H=cam.H, W=cam.W, args=(shs, col, sc, rot, cov, scale_modifier))
    return d


def ours_backward(scene, cam, fwd, grads, debug=False):
    from scorp_b200.diff_gaussian_rasterization import _C
    shs, col, sc, rot, cov, smod = fwd["args"]
    out = _C.rasterize_gaussians_backward(scene.bg, scene.means3D, fwd["radii"], col, sc, rot, smod, cov,
                                          cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0],
                                          grads[1], grads[2], shs, scene.sh_degree, cam.campos, fwd["geom"],
                                          fwd["num_rendered"], fwd["binning"], fwd["image"], fwd["alpha"], debug)
    keys = ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations")
    return dict(zip(keys, out))


def ours_intermediates(fwd):
    """Slice the product's private scratch (gsr_*_layout) into the reference's named arrays."""
    lib = _lib.lib()
    P, H, W, N = fwd["P"], fwd["H"], fwd["W"], fwd["num_rendered"]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    off = (C.c_size_t * 6)(); tot = C.c_size_t(0)
    lib.gsr_geom_layout(P, tiles, off, C.byref(tot))
    g = fwd["geom"]
    rec = g[off[0]: off[0] + 48 * P].view(torch.float32).view(P, 12)
    res = dict(means2D=rec[:, 0:2], conic_opacity=rec[:, 2:6], depths=rec[:, 6], rgb=rec[:, 8:11],
               cov3D=g[off[1]: off[1] + 24 * P].view(torch.float32).view(P, 6),
               clamped=g[off[2]: off[2] + P],
               tiles_touched=g[off[3]: off[3] + 4 * P].view(torch.int32))
    ioff = (C.c_size_t * 2)()
    lib.gsr_image_layout(W, H, ioff, C.byref(tot))
    im = fwd["image"]
    res["ranges"] = im[ioff[0]: ioff[0] + 8 * tiles].view(torch.int32).view(tiles, 2)
    res["n_contrib"] = im[ioff[1]: ioff[1] + 4 * H * W].view(torch.int32).view(H, W)
    boff = (C.c_size_t * 3)()
    lib.gsr_binning_layout(N, boff, C.byref(tot))
    b = fwd["binning"]
    if N > 0:
        res["point_list"] = b[boff[0]: boff[0] + 4 * N].view(torch.int32)
        res["point_list_keys"] = b[boff[1]: boff[1] + 8 * N].view(torch.int64)
    else:
        res["point_list"] = torch.zeros(0, dtype=torch.int32, device=g.device)
        res["point_list_keys"] = torch.zeros(0, dtype=torch.int64, device=g.device)
    return res


def ref_forward(scene, cam, colors_precomp=None, cov3D_precomp=None, scale_modifier=1.0):
    from oracle import ref_cuda
    shs = None if colors_precomp is not None else scene.shs
    sc = None if cov3D_precomp is not None else scene.scales
    rot = None if cov3D_precomp is not None else scene.rotations
    out = ref_cuda.dgr_forward(scene.bg, scene.means3D, colors_precomp, scene.opacities, sc, rot, scale_modifier,
                               cov3D_precomp, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H,
                               cam.W, shs, scene.sh_degree, cam.campos)
    out["args"] = (shs, colors_precomp, sc, rot, cov3D_precomp, scale_modifier)
    return out


def ref_backward(scene, cam, fwd, grads):
    from oracle import ref_cuda
    shs, col, sc, rot, cov, smod = fwd["args"]
    return ref_cuda.dgr_backward(fwd, scene.bg, scene.means3D, col, sc, rot, smod, cov, cam.viewmatrix,
                                 cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0], grads[1], grads[2], shs,
                                 scene.sh_degree, cam.campos)


def close(a, b, rtol=1e-4, atol_scale=1e-6):
    """rtol 1e-4 with atol = 1e-6 * max|b|. Returns (ok, max_violation_ratio)."""
    a = a.double().flatten(); b = b.double().flatten()
    atol = atol_scale * float(b.abs().max()) if b.numel() else 0.0
    err = (a - b).abs()
    tol = atol + rtol * b.abs()
    bad = err > tol
    return (not bool(bad.any())), float((err / tol.clamp_min(1e-300)).max()) if b.numel() else 0.0


def close_to_runs(a, runs, rtol=1e-4, atol_scale=1e-6):
    """`a` against SEVERAL runs of the same computation whose results differ only by the order of float atomics: within
    rtol of their mean, atol = atol_scale * max|mean| + 2 x their own largest run-to-run difference (what
    test_gpu_dgr_parity.py:_grad_ok does with the reference runs). Returns (ok, max_violation_ratio)."""
    st = torch.stack([r.double().flatten() for r in runs])
    mean = st.mean(0)
    if mean.numel() == 0:
        return True, 0.0
    spread = float((st.max(0).values - st.min(0).values).max()) if len(runs) > 1 else 0.0
    tol = atol_scale * float(mean.abs().max()) + 2.0 * spread + rtol * mean.abs()
    err = (a.double().flatten() - mean).abs()
    return (not bool((err > tol).any())), float((err / tol.clamp_min(1e-300)).max())


def make(name, device="cuda", **kw):
    return synth.make_scene(name, **kw).to(device)


# ---- 2DGS surfels ---------------------------------------------------------------------------------

def ours_surfel_forward(scene, cam, colors_precomp=None, transMat_precomp=None, scale_modifier=1.0, debug=False):
    from scorp_b200.diff_surfel_rasterization import _C
    shs = empty() if colors_precomp is not None else scene.shs
    col = colors_precomp if colors_precomp is not None else empty()
    sc = empty() if transMat_precomp is not None else scene.scales
    rot = empty() if transMat_precomp is not None else scene.rotations
    tm = transMat_precomp if transMat_precomp is not None else empty()
    out = _C.rasterize_gaussians(scene.bg, scene.means3D, col, scene.opacities, sc, rot, scale_modifier, tm,
                                 cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, shs,
                                 scene.sh_degree, cam.campos, False, debug)
    keys = ("num_rendered", "color", "others", "radii", "geom", "binning", "image")
    d = dict(zip(keys, out))
    d.update(P=scene.P, H=cam.H, W=cam.W, args=(shs, col, sc, rot, tm, scale_modifier))
    return d


def ours_surfel_backward(scene, cam, fwd, grads, debug=False):
    from scorp_b200.diff_surfel_rasterization import _C
    shs, col, sc, rot, tm, smod = fwd["args"]
    out = _C.rasterize_gaussians_backward(scene.bg, scene.means3D, fwd["radii"], col, sc, rot, smod, tm,
                                          cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0],
                                          grads[1], shs, scene.sh_degree, cam.campos, fwd["geom"],
                                          fwd["num_rendered"], fwd["binning"], fwd["image"], debug)
    keys = ("means2D", "colors", "opacity", "means3D", "transMat", "sh", "scales", "rotations")
    return dict(zip(keys, out))


def ours_surfel_intermediates(fwd):
    lib = _lib.lib()
    P, H, W, N = fwd["P"], fwd["H"], fwd["W"], fwd["num_rendered"]
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    off = (C.c_size_t * 6)(); tot = C.c_size_t(0)
    lib.gsr_surfel_geom_layout(P, tiles, off, C.byref(tot))
    g = fwd["geom"]
    rec = g[off[0]: off[0] + 96 * P].view(torch.float32).view(P, 24)
    res = dict(means2D=rec[:, 0:2], transMat=rec[:, 2:11], normal_opacity=rec[:, 11:15], rgb=rec[:, 15:18],
               depths=rec[:, 18], box=rec[:, 20:24],
               clamped=g[off[1]: off[1] + P],
               tiles_touched=g[off[2]: off[2] + 4 * P].view(torch.int32))
    ioff = (C.c_size_t * 4)()
    lib.gsr_surfel_image_layout(W, H, ioff, C.byref(tot))
    im = fwd["image"]
    stride = ioff[2] - ioff[1]
    plane = lambda k, dt: im[ioff[1] + k * stride: ioff[1] + k * stride + 4 * H * W].view(dt).view(H, W)
    res["ranges"] = im[ioff[0]: ioff[0] + 8 * tiles].view(torch.int32).view(tiles, 2)
    res["n_contrib"] = plane(0, torch.int32); res["median_contrib"] = plane(1, torch.int32)
    res["final_T"] = plane(2, torch.float32); res["M1"] = plane(3, torch.float32); res["M2"] = plane(4, torch.float32)
    boff = (C.c_size_t * 3)()
    lib.gsr_binning_layout(N, boff, C.byref(tot))
    b = fwd["binning"]
    if N > 0:
        res["point_list"] = b[boff[0]: boff[0] + 4 * N].view(torch.int32)
        res["point_list_keys"] = b[boff[1]: boff[1] + 8 * N].view(torch.int64)
    else:
        res["point_list"] = torch.zeros(0, dtype=torch.int32, device=g.device)
        res["point_list_keys"] = torch.zeros(0, dtype=torch.int64, device=g.device)
    return res


def ref_surfel_forward(scene, cam, colors_precomp=None, transMat_precomp=None, scale_modifier=1.0):
    from oracle import ref_cuda
    shs = None if colors_precomp is not None else scene.shs
    sc = None if transMat_precomp is not None else scene.scales
    rot = None if transMat_precomp is not None else scene.rotations
    out = ref_cuda.dsr_forward(scene.bg, scene.means3D, colors_precomp, scene.opacities, sc, rot, scale_modifier,
                               transMat_precomp, cam.viewmatrix, cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H,
                               cam.W, shs, scene.sh_degree, cam.campos)
    out["args"] = (shs, colors_precomp, sc, rot, transMat_precomp, scale_modifier)
    return out


def ref_surfel_backward(scene, cam, fwd, grads):
    from oracle import ref_cuda
    shs, col, sc, rot, tm, smod = fwd["args"]
    return ref_cuda.dsr_backward(fwd, scene.bg, scene.means3D, col, sc, rot, smod, tm, cam.viewmatrix,
                                 cam.projmatrix, cam.tanfovx, cam.tanfovy, grads[0], grads[1], shs,
                                 scene.sh_degree, cam.campos)


def sh_basis_torch(D, dirs):
    """Real SH basis b_k(dir) for unit directions dirs[N,3] -> [N,(D+1)^2] (same constants as
    diff-gaussian-rasterization/cuda_rasterizer/auxiliary.h:20-36); test-side restatement used to check the
    factored multi-GPU exchange (scorp_b200/multiview.py:FactoredExchange)."""
    x, y, z = dirs[:, 0], dirs[:, 1], dirs[:, 2]
    C1 = 0.4886025119029199
    C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
    C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
          1.445305721320277, -0.5900435899266435)
    b = [torch.full_like(x, 0.28209479177387814)]
    if D > 0:
        b += [-C1 * y, C1 * z, -C1 * x]
    if D > 1:
        xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
        b += [C2[0] * xy, C2[1] * yz, C2[2] * (2 * zz - xx - yy), C2[3] * xz, C2[4] * (xx - yy)]
    if D > 2:
        b += [C3[0] * y * (3 * xx - yy), C3[1] * xy * z, C3[2] * y * (4 * zz - xx - yy),
              C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy), C3[5] * z * (xx - yy),
              C3[6] * x * (xx - 3 * yy)]
    return torch.stack(b, dim=1)


def sh_grad_from_views_torch(P, M, D, n_views, means3D, campos_all, drgb_all, dsh_out):
    """sum_v basis(normalize(mean - campos_v)) (x) drgb_v, written into dsh_out[P*M*3] (float64 accumulation)."""
    acc = torch.zeros((P, M, 3), dtype=torch.float64, device=means3D.device)
    for v in range(n_views):
        d = means3D.double() - campos_all[v].double()
        d = d / d.norm(dim=1, keepdim=True)
        b = sh_basis_torch(D, d)                                   # [P, used]
        acc[:, : b.shape[1], :] += b[:, :, None] * drgb_all[v].double().view(P, 1, 3)
    dsh_out.view(P, M, 3).copy_(acc.to(dsh_out.dtype))
