"""Diagnostic (not a test): product vs reference CUDA on the GPU box, prints every mismatch statistic."""
import sys, time, json
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT)); sys.path.insert(0, str(ROOT / "tests"))
import torch
import helpers as h
from scorp_b200 import synth

def stat(name, a, b, exact):
    if exact:
        if a.dtype.is_floating_point:
            a = a.contiguous().view(torch.int32); b = b.contiguous().view(torch.int32)
        n = int((a != b).sum())
        print(f"  {name:18s} exact-mismatch {n} / {a.numel()}")
    else:
        ok, worst = h.close(a, b)
        print(f"  {name:18s} close={ok} worst={worst:.3g} maxabs_ref={float(b.abs().max()):.3g} maxdiff={float((a-b).abs().max()):.3g}")

def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(n):
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    ts.sort(); return ts[len(ts)//2]

def run(name, **kw):
    sc = h.make(name, **kw); cam = sc.cameras[0]
    print(f"== {name} P={sc.P} D={sc.sh_degree} {cam.W}x{cam.H} {kw}")
    fo = h.ours_forward(sc, cam); fr = h.ref_forward(sc, cam)
    print("  num_rendered ours", fo["num_rendered"], "ref", fr["num_rendered"])
    io = h.ours_intermediates(fo)
    from oracle import ref_cuda
    ir = ref_cuda.dgr_intermediates(fr)
    stat("radii", fo["radii"], fr["radii"], True)
    vis = fr["radii"] > 0
    for k in ("means2D", "depths", "conic_opacity", "rgb", "cov3D", "tiles_touched"):
        stat(k, io[k][vis], ir[k][vis], True)
    if fo["num_rendered"] == fr["num_rendered"]:
        stat("point_list", io["point_list"], ir["point_list"], True)
        stat("point_list_keys", io["point_list_keys"], ir["point_list_keys"], True)
    stat("ranges", io["ranges"], ir["ranges"], True)
    stat("n_contrib", io["n_contrib"], ir["n_contrib"], True)
    for k in ("color", "depth", "alpha"):
        stat(k + "(bits)", fo[k], fr[k], True); stat(k, fo[k], fr[k], False)
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 1)]
    go = h.ours_backward(sc, cam, fo, grads); gr = h.ref_backward(sc, cam, fr, grads)
    gr2 = h.ref_backward(sc, cam, fr, grads)
    for k in ("means2D", "colors", "opacity", "means3D", "cov3D", "sh", "scales", "rotations"):
        stat("d" + k, go[k].reshape(-1), gr[k].reshape(-1), False)
        stat("d" + k + "(ref-ref)", gr2[k].reshape(-1), gr[k].reshape(-1), False)
    t_of = timeit(lambda: h.ours_forward(sc, cam)); t_rf = timeit(lambda: h.ref_forward(sc, cam))
    t_ob = timeit(lambda: h.ours_backward(sc, cam, fo, grads)); t_rb = timeit(lambda: h.ref_backward(sc, cam, fr, grads))
    print(f"  ms fwd ours {t_of:.3f} ref {t_rf:.3f} | bwd ours {t_ob:.3f} ref {t_rb:.3f}")

if __name__ == "__main__":
    run("C1", P=5000, W=200, H=136)
    run("C1")
    run("C2", P=100000)
    run("C2")
