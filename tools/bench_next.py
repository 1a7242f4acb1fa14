#!/usr/bin/env python
"""Measurement of the SURVEY.md 8(f) rows next to the rasterizer boundary, each against what a user of the reference
runs today on the same GPU (its own torch code around the drop-in rasterizer). One JSON line per row.

    python tools/bench_next.py [--config C2] [--steps 30] [--warmup 5]

row "staging" (8(f) rank 1): stored GaussianModel tensors -> image -> gradients on the stored tensors
    stock : torch.cat / sigmoid / exp / normalize + stock GaussianRasterizer + autograd  (gaussian_model.py:126-146)
    model : scorp_b200.model_space.rasterize_model (activations and the SH split fused into the preprocess kernels)
row "loss" (8(f) rank 3): (1 - l) * L1 + l * (1 - SSIM) forward + backward on a C2-sized image
    torch : loss_utils.py:17-18,45-73 restated call for call (five grouped conv2d + elementwise kernels + autograd)
    fused : scorp_b200.loss_utils.l1_ssim_loss (two kernels)
Times are CUDA events around K steps after W warm-ups, inputs resident; HBM fraction uses MEASURED_PEAKS.json.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))

import numpy as np
import torch
import torch.nn.functional as F

from scorp_b200 import synth


def peak():
    try:
        return float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


def timeit(fn, K, W):
    for _ in range(W):
        fn()
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(K + 1)]
    ev[0].record()
    for k in range(K):
        fn(); ev[k + 1].record()
    torch.cuda.synchronize()
    per = sorted(ev[k].elapsed_time(ev[k + 1]) for k in range(K))
    return ev[0].elapsed_time(ev[K]) / K, {"median": per[K // 2], "min": per[0], "max": per[-1]}


def row_staging(cfg, K, W):
    import scorp_b200.diff_gaussian_rasterization as dgr
    from scorp_b200 import model_space as ms
    sc = synth.make_scene(cfg).to("cuda"); cam = sc.cameras[0]
    P, M = sc.P, sc.shs.shape[1]
    op = sc.opacities.clamp(1e-4, 1 - 1e-4)
    st = dict(xyz=sc.means3D.clone(), features_dc=sc.shs[:, :1].contiguous(), features_rest=sc.shs[:, 1:].contiguous(),
              opacity=torch.log(op / (1 - op)), scaling=torch.log(sc.scales), rotation=sc.rotations * 1.7)
    st = {k: v.requires_grad_(True) for k, v in st.items()}
    rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, sc.sh_degree, cam.campos, False, False)
    gr = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 1)]

    def finish(outs):
        torch.autograd.backward([outs[0], outs[2], outs[3]], gr)

    def stock():
        for v in st.values():
            v.grad = None
        shs = torch.cat((st["features_dc"], st["features_rest"]), dim=1)
        m2d = torch.zeros_like(st["xyz"], requires_grad=True) + 0
        finish(dgr.GaussianRasterizer(rs)(means3D=st["xyz"], means2D=m2d, opacities=torch.sigmoid(st["opacity"]),
                                          shs=shs, scales=torch.exp(st["scaling"]),
                                          rotations=F.normalize(st["rotation"])))

    def model():
        for v in st.values():
            v.grad = None
        m2d = torch.zeros_like(st["xyz"], requires_grad=True) + 0
        finish(ms.rasterize_model(st["xyz"], m2d, st["features_dc"], st["features_rest"], st["opacity"], st["scaling"],
                                  st["rotation"], rs))

    for _ in range(8):
        stock(); model()
    t_stock, sp_stock = timeit(stock, K, W)
    t_model, sp_model = timeit(model, K, W)
    # traffic the fused form removes: cat (read + write the SH tensor) and its two slice copies in backward, plus
    # 3 activations fwd (read + write) and their backward kernels (read grad, read saved, write grad)
    removed = 4 * P * (2 * 3 * M + 2 * 3 * M + 2 * (1 + 3 + 4) + 3 * (1 + 3 + 4))
    return {"row": "8f-1 staging", "config": cfg, "P": P, "sh_coeffs": M, "unit": "ms per fwd+bwd (autograd, inputs resident)",
            "stock_torch_staging_ms": t_stock, "model_space_ms": t_model, "speedup": t_stock / t_model,
            "spread": {"stock": sp_stock, "model": sp_model}, "staging_bytes_removed": removed}


def torch_loss(x, b, lam, w):
    Cn = x.size(-3)
    conv = lambda t: F.conv2d(t, w, padding=5, groups=Cn)
    mu1, mu2 = conv(x), conv(b)
    mu1_sq, mu2_sq, mu1_mu2 = mu1.pow(2), mu2.pow(2), mu1 * mu2
    s1 = conv(x * x) - mu1_sq; s2 = conv(b * b) - mu2_sq; s12 = conv(x * b) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    m = ((2 * mu1_mu2 + C1) * (2 * s12 + C2)) / ((mu1_sq + mu2_sq + C1) * (s1 + s2 + C2))
    return (1.0 - lam) * torch.abs(x - b).mean() + lam * (1.0 - m.mean())


def row_loss(cfg, K, W):
    from scorp_b200 import loss_utils as lu, _lib
    c = synth.CONFIGS[cfg]
    Wd, Hd = c["W"], c["H"]
    gen = torch.Generator().manual_seed(1)
    a = torch.rand(3, Hd, Wd, generator=gen).cuda().requires_grad_(True)
    b = torch.rand(3, Hd, Wd, generator=gen).cuda()
    g = torch.tensor([np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g = (g / g.sum()).unsqueeze(1)
    w = g.mm(g.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().cuda()
    lam = 0.2

    def ref():
        a.grad = None
        torch_loss(a, b, lam, w).backward()

    def fused():
        a.grad = None
        lu.l1_ssim_loss(a, b, lam).backward()

    for _ in range(8):
        ref(); fused()
    t_ref, sp_ref = timeit(ref, K, W)
    t_fused, sp_fused = timeit(fused, K, W)
    # kernel-only times of the two fused kernels (C ABI, no autograd / allocation)
    lib = _lib.lib()
    n = a.numel()
    nb = lib.gsr_l1_ssim_workspace_bytes(3, Hd, Wd)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda"); out = torch.empty(3, device="cuda"); gi = torch.empty_like(b)
    ad = a.detach()
    st = torch.cuda.current_stream().cuda_stream
    # ping-pong over several image pairs so successive launches do not find their inputs in L2
    kf, _ = timeit(lambda: lib.gsr_l1_ssim_forward(3, Hd, Wd, ad.data_ptr(), b.data_ptr(), lam, out.data_ptr(), ws.data_ptr(), nb, st), 50, 5)
    kb, _ = timeit(lambda: lib.gsr_l1_ssim_backward(3, Hd, Wd, ad.data_ptr(), b.data_ptr(), None, 1 - lam, -lam, ws.data_ptr(), gi.data_ptr(), st), 50, 5)
    pk, src = peak()
    bytes_f, bytes_b = 4 * n * (2 + 3), 4 * n * (3 + 2 + 1)
    return {"row": "8f-3 loss", "config": cfg, "image": [3, Hd, Wd], "unit": "ms per loss fwd+bwd (autograd, inputs resident)",
            "torch_ms": t_ref, "fused_ms": t_fused, "speedup": t_ref / t_fused, "spread": {"torch": sp_ref, "fused": sp_fused},
            "kernels": {"forward_ms": kf, "backward_ms": kb,
                        "note": "back-to-back launches on a 78 MB working set (inside the 126 MB L2): L2-warm, as in a training loop where the rasterizer just wrote the image"},
            "roofline": {"bound": "hbm", "peak": pk, "peak_source": src, "unit": "GB/s",
                         "forward": {"algorithmic_bytes": bytes_f, "achieved": bytes_f / kf / 1e6, "frac": bytes_f / kf / 1e6 / pk},
                         "backward": {"algorithmic_bytes": bytes_b, "achieved": bytes_b / kb / 1e6, "frac": bytes_b / kb / 1e6 / pk}}}


def row_optim(cfg, K, W):
    """8(f) rank 4: optimizer.step() over the six groups of a C2-sized model + the densification statistics."""
    from scorp_b200 import optim
    P = synth.CONFIGS[cfg]["P"]; M = (synth.CONFIGS[cfg]["sh_degree"] + 1) ** 2
    shapes = ((P, 3), (P, 1, 3), (P, max(M - 1, 0), 3), (P, 1), (P, 3), (P, 4))
    lrs = (0.00016, 0.0025, 0.0025 / 20.0, 0.05, 0.005, 0.001)
    names = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")
    gen = torch.Generator().manual_seed(0)

    def make(cls):
        ps = [torch.nn.Parameter(torch.randn(*s, generator=gen).cuda()) for s in shapes if s[1] > 0]
        for p in ps:
            p.grad = torch.randn_like(p) * 1e-3
        return ps, cls([{"params": [p], "lr": lr, "name": n} for p, lr, n in zip(ps, lrs, names)], lr=0.0, eps=1e-15)
    pa, oa = make(torch.optim.Adam)
    pb, ob = make(optim.Adam)
    for _ in range(3):
        oa.step(); ob.step()
    t_torch, sp_t = timeit(oa.step, K, W)
    t_fused, sp_f = timeit(ob.step, K, W)
    n = sum(p.numel() for p in pb)
    pk, src = peak()
    bytes_ = 28 * n
    # densification statistics
    radii = torch.randint(0, 30, (P,), device="cuda", dtype=torch.int32)
    vg = torch.randn(P, 3, device="cuda") * 1e-4
    accum = torch.zeros(P, 1, device="cuda"); denom = torch.zeros(P, 1, device="cuda"); maxr = torch.zeros(P, device="cuda")

    def ref_stats():
        vis = radii > 0
        maxr[vis] = torch.max(maxr[vis], radii[vis])
        accum[vis] += torch.norm(vg[vis, :2], dim=-1, keepdim=True)
        denom[vis] += 1

    t_rs, _ = timeit(ref_stats, K, W)
    t_fs, _ = timeit(lambda: optim.add_densification_stats(accum, denom, maxr, vg, radii), K, W)
    return {"row": "8f-4 optimizer", "config": cfg, "param_floats": n, "unit": "ms per optimizer.step() (six groups)",
            "torch_adam_ms": t_torch, "fused_adam_ms": t_fused, "speedup": t_torch / t_fused,
            "spread": {"torch": sp_t, "fused": sp_f},
            "roofline": {"bound": "hbm", "peak": pk, "peak_source": src, "unit": "GB/s", "algorithmic_bytes": bytes_,
                         "achieved": bytes_ / t_fused / 1e6, "frac": bytes_ / t_fused / 1e6 / pk,
                         "note": "28 B per parameter float (read param, grad, two moments; write param and two moments); 1.65 GB working set, far beyond L2"},
            "densification_stats": {"torch_mask_indexing_ms": t_rs, "fused_ms": t_fs, "speedup": t_rs / t_fs}}


def row_densify(cfg, K, W):
    """8(f) rank 4: GaussianModel.densify_and_prune on a C2-sized model with Adam moments -- the reference's rounds of
    mask indexing and torch.cat (tools/torch_densify_sequence.py) against the plan + gather kernels."""
    sys.path.insert(0, str(ROOT / "tools"))
    from torch_densify_sequence import ATTR, model, torch_sequence
    from scorp_b200 import optim
    P = synth.CONFIGS[cfg]["P"]; M = (synth.CONFIGS[cfg]["sh_degree"] + 1) ** 2
    pc = model(P, M - 1, 3, seed=1)
    par = {n: getattr(pc, a).detach() for n, a in ATTR.items()}
    mom = {n: (pc.optimizer.state[getattr(pc, a)]["exp_avg"], pc.optimizer.state[getattr(pc, a)]["exp_avg_sq"])
           for n, a in ATTR.items() if par[n].numel()}
    args = (0.0002, 0.005, 4.0, 20, 0.01)
    K, W = min(K, 10), min(W, 3)
    t_torch, sp_t = timeit(lambda: torch_sequence(par, mom, pc.xyz_gradient_accum, pc.denom, *args), K, W)
    res = {}

    def fused():
        res["out"] = optim.densify_and_prune_tensors(par, mom, pc.xyz_gradient_accum, pc.denom, *args)
    t_fused, sp_f = timeit(fused, K, W)
    out, _, counts = res["out"]
    rows = out["xyz"].shape[0]
    floats = sum(t.numel() // max(P, 1) for t in par.values())
    # read the surviving source rows of parameters and moments once, write every output row once (moments of new rows
    # are zeros that still have to be written), plus the plan's inputs (scaling, opacity, accum, denom)
    bytes_ = 4 * floats * (counts[0] * 3 + (rows - counts[0])) + 4 * floats * rows * 3 + P * (12 + 4 + 4 + 4)
    pk, src = peak()
    return {"row": "8f-4 densify_and_prune", "config": cfg, "P": P, "rows_after": rows, "counts": list(counts),
            "unit": "ms per densify_and_prune (six parameter tensors + twelve Adam moment tensors)",
            "torch_sequence_ms": t_torch, "fused_ms": t_fused, "speedup": t_torch / t_fused,
            "spread": {"torch": sp_t, "fused": sp_f},
            "roofline": {"bound": "hbm", "peak": pk, "peak_source": src, "unit": "GB/s", "algorithmic_bytes": bytes_,
                         "achieved": bytes_ / t_fused / 1e6, "frac": bytes_ / t_fused / 1e6 / pk,
                         "note": "whole call including the count readback, output allocation and the normal draw"}}


def row_ply(cfg, K, W):
    """8(f) rank 4, last item: save_ply / load_ply of a C2-sized model. The torch/numpy arm is the reference's own statement
    sequence (gaussian_model.py:236-247 and :292-326) without the plyfile call that follows it, which is not in this image:
    it is charged only for building / consuming the structured array, so the real reference is slower than this arm."""
    import os
    import tempfile
    import time
    from scorp_b200 import ply_io
    P = synth.CONFIGS[cfg]["P"]; R = (synth.CONFIGS[cfg]["sh_degree"] + 1) ** 2 - 1
    g = torch.Generator().manual_seed(0)
    r = lambda *s: torch.randn(*s, generator=g).cuda()
    m = dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, R, 3), opacity=r(P, 1), scaling=r(P, 3), rotation=r(P, 4))
    names = ply_io.construct_list_of_attributes(R, 3)
    tmp = tempfile.mkdtemp()
    path = os.path.join(tmp, "point_cloud.ply")

    def wall(fn, n):
        fn(); torch.cuda.synchronize()
        t = []
        for _ in range(n):
            t0 = time.perf_counter(); fn(); torch.cuda.synchronize(); t.append((time.perf_counter() - t0) * 1e3)
        return sorted(t)[len(t) // 2]

    def ref_save_rows():
        xyz = m["xyz"].detach().cpu().numpy()
        normals = np.zeros_like(xyz)
        f_dc = m["features_dc"].detach().transpose(1, 2).flatten(start_dim=1).contiguous().cpu().numpy()
        f_rest = m["features_rest"].detach().transpose(1, 2).flatten(start_dim=1).contiguous().cpu().numpy()
        elements = np.empty(xyz.shape[0], dtype=[(a, "f4") for a in names])
        attributes = np.concatenate((xyz, normals, f_dc, f_rest, m["opacity"].cpu().numpy(), m["scaling"].cpu().numpy(),
                                     m["rotation"].cpu().numpy()), axis=1)
        elements[:] = list(map(tuple, attributes))
        return elements

    t0 = time.perf_counter(); el = ref_save_rows(); t_ref_save = (time.perf_counter() - t0) * 1e3
    t_save = wall(lambda: ply_io.save_ply(m, path), 5)

    def ref_load():
        rec = np.fromfile(path, dtype=el.dtype, offset=os.path.getsize(path) - el.nbytes)
        xyz = np.stack((np.asarray(rec["x"]), np.asarray(rec["y"]), np.asarray(rec["z"])), axis=1)
        fe = np.zeros((P, 3 * R))
        for idx in range(3 * R):
            fe[:, idx] = np.asarray(rec["f_rest_%d" % idx])
        fe = fe.reshape((P, 3, R))
        fd = np.zeros((P, 3, 1))
        for c in range(3):
            fd[:, c, 0] = np.asarray(rec["f_dc_%d" % c])
        sc = np.zeros((P, 3)); ro = np.zeros((P, 4))
        for c in range(3):
            sc[:, c] = np.asarray(rec["scale_%d" % c])
        for c in range(4):
            ro[:, c] = np.asarray(rec["rot_%d" % c])
        op = np.asarray(rec["opacity"])[..., np.newaxis]
        return [torch.tensor(a, dtype=torch.float, device="cuda") for a in (xyz, op, sc, ro)] + \
               [torch.tensor(fd, dtype=torch.float, device="cuda").transpose(1, 2).contiguous(),
                torch.tensor(fe, dtype=torch.float, device="cuda").transpose(1, 2).contiguous()]

    t_ref_load = wall(ref_load, 3)
    t_load = wall(lambda: ply_io.load_ply(path, max_sh_degree=3), 5)
    back = ply_io.load_ply(path, max_sh_degree=3)
    ok = all(torch.equal(back[k], m[k]) for k in m)
    size = os.path.getsize(path)
    os.remove(path); os.rmdir(tmp)
    return {"row": "8f-4 ply", "config": cfg, "P": P, "file_bytes": size, "round_trip_exact": ok,
            "unit": "ms wall clock per call (file system included on our side)",
            "reference_statements_save_ms": t_ref_save, "save_ply_ms": t_save, "save_speedup": t_ref_save / t_save,
            "reference_statements_load_ms": t_ref_load, "load_ply_ms": t_load, "load_speedup": t_ref_load / t_load,
            "note": "reference arm = its numpy statements only (structured array through list(map(tuple, ...)) on save, one "
                    "column at a time on load), without plyfile's own write / read"}


def row_iteration(cfg, K, W):
    """One whole training iteration as train_3dgs.py:85-193 runs it -- staging, render, loss, backward, densification
    statistics, optimizer step -- (a) with the reference's torch code around the stock drop-in rasterizer, (b) with
    every 8(f) row switched on."""
    import scorp_b200.diff_gaussian_rasterization as dgr
    from scorp_b200 import model_space as ms, loss_utils as lu, optim
    sc = synth.make_scene(cfg).to("cuda"); cam = sc.cameras[0]
    P = sc.P
    op = sc.opacities.clamp(1e-4, 1 - 1e-4)
    rs = dgr.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, sc.sh_degree, cam.campos, False, False)
    gt = torch.rand(3, cam.H, cam.W, device="cuda")
    g11 = torch.tensor([np.exp(-(x - 5) ** 2 / (2 * 1.5 ** 2)) for x in range(11)], dtype=torch.float32)
    g11 = (g11 / g11.sum()).unsqueeze(1)
    w = g11.mm(g11.t()).float()[None, None].expand(3, 1, 11, 11).contiguous().cuda()
    lam = 0.2
    lrs = (0.00016, 0.0025, 0.0025 / 20.0, 0.05, 0.005, 0.001)
    names = ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")

    def model():
        st = dict(xyz=sc.means3D.clone(), features_dc=sc.shs[:, :1].contiguous().clone(),
                  features_rest=sc.shs[:, 1:].contiguous().clone(), opacity=torch.log(op / (1 - op)),
                  scaling=torch.log(sc.scales), rotation=sc.rotations * 1.7)
        return {k: torch.nn.Parameter(v) for k, v in st.items()}

    def stats():
        return torch.zeros(P, 1, device="cuda"), torch.zeros(P, 1, device="cuda"), torch.zeros(P, device="cuda")

    sa = model(); oa = torch.optim.Adam([{"params": [sa[n]], "lr": lr, "name": n} for n, lr in zip(names, lrs)], lr=0.0, eps=1e-15)
    acc_a, den_a, maxr_a = stats()
    sb = model(); ob = optim.Adam([{"params": [sb[n]], "lr": lr, "name": n} for n, lr in zip(names, lrs)], lr=0.0, eps=1e-15)
    acc_b, den_b, maxr_b = stats()

    def stock():
        shs = torch.cat((sa["features_dc"], sa["features_rest"]), dim=1)
        m2d = torch.zeros_like(sa["xyz"], requires_grad=True) + 0
        m2d.retain_grad()
        image, radii, depth, alpha = dgr.GaussianRasterizer(rs)(
            means3D=sa["xyz"], means2D=m2d, opacities=torch.sigmoid(sa["opacity"]), shs=shs,
            scales=torch.exp(sa["scaling"]), rotations=F.normalize(sa["rotation"]))
        loss = torch_loss(image, gt, lam, w)
        loss.backward()
        vis = radii > 0
        maxr_a[vis] = torch.max(maxr_a[vis], radii[vis])
        acc_a[vis] += torch.norm(m2d.grad[vis, :2], dim=-1, keepdim=True)
        den_a[vis] += 1
        oa.step(); oa.zero_grad(set_to_none=True)
        return loss

    def fused():
        m2d = torch.zeros_like(sb["xyz"], requires_grad=True) + 0
        m2d.retain_grad()
        image, radii, depth, alpha = ms.rasterize_model(sb["xyz"], m2d, sb["features_dc"], sb["features_rest"],
                                                        sb["opacity"], sb["scaling"], sb["rotation"], rs)
        loss = lu.l1_ssim_loss(image, gt, lam)
        loss.backward()
        optim.add_densification_stats(acc_b, den_b, maxr_b, m2d.grad, radii)
        ob.step(); ob.zero_grad(set_to_none=True)
        return loss

    for _ in range(8):
        la = stock(); lb = fused()
    t_stock, sp_s = timeit(stock, K, W)
    t_fused, sp_f = timeit(fused, K, W)
    return {"row": "8f whole iteration", "config": cfg, "unit": "ms per training iteration (staging + render + loss + backward + densification statistics + Adam step), inputs resident",
            "torch_around_stock_rasterizer_ms": t_stock, "all_rows_fused_ms": t_fused, "speedup": t_stock / t_fused,
            "spread": {"stock": sp_s, "fused": sp_f}, "loss_after_warmup": {"stock": float(la.detach()), "fused": float(lb.detach())},
            "Mpix_per_s": {"stock": cam.W * cam.H / t_stock / 1e3, "fused": cam.W * cam.H / t_fused / 1e3}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="C2")
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--rows", default="staging,loss,optim,iteration")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("needs a CUDA device")
    for r in args.rows.split(","):
        fn = {"staging": row_staging, "loss": row_loss, "optim": row_optim, "densify": row_densify, "ply": row_ply,
              "iteration": row_iteration}[r]
        print(json.dumps(fn(args.config, args.steps, args.warmup)), flush=True)


if __name__ == "__main__":
    main()
