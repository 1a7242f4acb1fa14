#!/usr/bin/env python
"""GPU diagnostic: surfel path and knn of libgsr_b200.so vs the reference CUDA build (oracle/_ref).
Prints mismatch counts per intermediate and gradient error statistics; exits 0 always."""
import sys
import time
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import helpers as h  # noqa: E402
from scorp_b200 import synth  # noqa: E402
from oracle import ref_cuda  # noqa: E402


def bits(t):
    return t.contiguous().view(torch.int32) if t.dtype.is_floating_point else t


def cmp_exact(name, a, b):
    a, b = bits(a), bits(b)
    n = int((a != b).sum())
    print(f"  {name:18s} mismatches {n:8d} / {a.numel()}")
    return n


def cmp_close(name, a, b, spread=None):
    a = a.double().flatten(); b = b.double().flatten()
    err = (a - b).abs()
    scale = float(b.abs().max())
    tol = 1e-6 * scale + 1e-4 * b.abs() + (2 * spread if spread is not None else 0)
    bad = int((err > tol).sum())
    print(f"  {name:18s} max|err| {float(err.max()):.3e}  max|ref| {scale:.3e}  beyond-tol {bad} / {a.numel()}"
          + (f"  ref-spread {spread:.2e}" if spread is not None else ""))
    return bad


def run_case(name, cfg, kw, bwd=True, views=(0,)):
    sc = h.make(cfg, **kw)
    for v in views:
        cam = sc.cameras[v]
        print(f"== {name} view {v}: P={sc.P} {cam.W}x{cam.H}")
        fo = h.ours_surfel_forward(sc, cam); fr = h.ref_surfel_forward(sc, cam)
        print("  num_rendered", fo["num_rendered"], fr["num_rendered"])
        io, ir = h.ours_surfel_intermediates(fo), ref_cuda.dsr_intermediates(fr)
        cmp_exact("radii", fo["radii"], fr["radii"])
        vis = (fr["radii"] > 0) & (fo["radii"] > 0)
        for k in ("depths", "transMat", "means2D", "normal_opacity", "rgb", "tiles_touched"):
            cmp_exact(k, io[k][vis], ir[k][vis])
        if fo["num_rendered"] == fr["num_rendered"]:
            cmp_exact("point_list_keys", io["point_list_keys"], ir["point_list_keys"])
            cmp_exact("point_list", io["point_list"], ir["point_list"])
            cmp_exact("ranges", io["ranges"], ir["ranges"])
        cmp_exact("n_contrib", io["n_contrib"], ir["n_contrib"][0])
        cmp_exact("median_contrib", io["median_contrib"], ir["n_contrib"][1])
        cmp_exact("final_T", io["final_T"], ir["accum"][0])
        cmp_exact("M1", io["M1"], ir["accum"][1])
        cmp_exact("M2", io["M2"], ir["accum"][2])
        cmp_exact("color", fo["color"], fr["color"])
        cmp_exact("others", fo["others"], fr["others"])
        cmp_close("color~", fo["color"], fr["color"])
        cmp_close("others~", fo["others"], fr["others"])
        if not bwd:
            continue
        grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 5, channels=(3, 7))]
        go = h.ours_surfel_backward(sc, cam, fo, grads)
        runs = [h.ref_surfel_backward(sc, cam, fr, grads) for _ in range(4)]
        for k in ("means2D", "colors", "opacity", "means3D", "transMat", "sh", "scales", "rotations"):
            st = torch.stack([r[k].double().reshape(-1) for r in runs])
            spread = float((st.max(0).values - st.min(0).values).max())
            cmp_close("dL_d" + k, go[k], st.mean(0), spread)


def knn_cases():
    g = torch.Generator().manual_seed(11)
    from scorp_b200.simple_knn._C import distCUDA2
    cases = {
        "uniform_100k": torch.rand(100_000, 3, generator=g) * 8 - 4,
        "clustered_200k": torch.cat([torch.randn(50_000, 3, generator=g) * s + c for s, c in
                                     ((0.05, 0.0), (0.5, 3.0), (2.0, -5.0), (0.001, 1.0))]),
        "planar_50k": torch.cat([torch.rand(50_000, 2, generator=g), torch.zeros(50_000, 1)], 1),
        "dups_20k": torch.rand(5_000, 3, generator=g).repeat(4, 1),
        "tiny_5": torch.rand(5, 3, generator=g),
        "far_offset_100k": torch.rand(100_000, 3, generator=g) * 2 + 1000.0,
        "uniform_1M": torch.rand(1_000_000, 3, generator=g) * 8 - 4,
    }
    for name, pts in cases.items():
        pts = pts.cuda().contiguous()
        ours = distCUDA2(pts); torch.cuda.synchronize()
        ref = ref_cuda.knn_dist2(pts)
        t0 = time.perf_counter(); distCUDA2(pts); torch.cuda.synchronize(); t_o = time.perf_counter() - t0
        t0 = time.perf_counter(); ref_cuda.knn_dist2(pts); t_r = time.perf_counter() - t0
        n = int((ours.view(torch.int32) != ref.view(torch.int32)).sum())
        print(f"knn {name:16s} P={pts.shape[0]:8d} bit-mismatches {n:6d} max|d| {float((ours - ref).abs().max()):.3e} "
              f"ours {t_o * 1e3:.2f} ms ref {t_r * 1e3:.2f} ms")


if __name__ == "__main__":
    torch.manual_seed(0)
    knn_cases()
    run_case("tiny", "C4", dict(P=5000, W=200, H=136))
    run_case("sh0_bg", "C4", dict(P=30_000, W=320, H=200, sh_degree=0, random_bg=True))
    run_case("C4_100k", "C4", dict(P=100_000))
    run_case("orbit", "C4", dict(P=50_000, W=400, H=260, n_views=3), views=(1, 2))
    run_case("C4_full", "C4", dict(), bwd=True)
