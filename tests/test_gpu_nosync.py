"""Forward without the host round trip (gsr_rasterize_forward_nosync / gsr_surfel_forward_nosync, scorp_b200/capacity.py).

The reference synchronises in the middle of every forward to read num_rendered (DGR / DSR
cuda_rasterizer/rasterizer_impl.cu:278-286); the no-sync variants size the binning buffer from a caller-supplied
capacity. With enough capacity every output and the sorted lists must be bit-identical to the stock forward (whose parity
with the reference CUDA is established by test_gpu_dgr_parity.py / test_gpu_dsr_parity.py), in all three size classes of
the per-tile sort; beyond the capacity the overflow flag must be raised and nothing binned. Also: the public API under
capacity.enable(), and forward + backward captured in a CUDA graph."""
import pytest
import torch

import helpers as h
from scorp_b200 import capacity, synth

pytestmark = pytest.mark.gpu


def _nosync_forward(sc, cam, surfel, cap):
    if surfel:
        from scorp_b200.diff_surfel_rasterization import _C
    else:
        from scorp_b200.diff_gaussian_rasterization import _C
    E = h.empty()
    status = torch.full((4,), -1, dtype=torch.int32).pin_memory()
    out = _C.rasterize_gaussians_nosync(sc.bg, sc.means3D, E, sc.opacities, sc.scales, sc.rotations, 1.0, E, cam.viewmatrix,
                                        cam.projmatrix, cam.tanfovx, cam.tanfovy, cam.H, cam.W, sc.shs, sc.sh_degree,
                                        cam.campos, False, False, cap, status)
    keys = (("num_rendered", "color", "others", "radii", "geom", "binning", "image") if surfel else
            ("num_rendered", "color", "depth", "alpha", "radii", "geom", "binning", "image"))
    d = dict(zip(keys, out))
    d.update(P=sc.P, H=cam.H, W=cam.W, args=(sc.shs, E, sc.scales, sc.rotations, E, 1.0), status=status)
    return d


CASES = [
    ("3dgs_small_lists", False, "C2", dict(P=100_000)),
    ("3dgs_mid_lists", False, "C1", dict(P=60_000, W=96, H=64)),
    ("3dgs_long_lists", False, "C1", dict(P=200_000, W=96, H=64)),
    ("2dgs", True, "C4", dict(P=100_000)),
    ("2dgs_mid_lists", True, "C4", dict(P=150_000, W=96, H=64, sh_degree=1)),
]


@pytest.mark.parametrize("name,surfel,cfg,kw", CASES, ids=[c[0] for c in CASES])
def test_nosync_forward_is_bit_identical_to_the_stock_forward(name, surfel, cfg, kw):
    sc = h.make(cfg, **kw); cam = sc.cameras[0]
    fwd_f, bwd_f, inter = ((h.ours_surfel_forward, h.ours_surfel_backward, h.ours_surfel_intermediates) if surfel else
                           (h.ours_forward, h.ours_backward, h.ours_intermediates))
    fs = fwd_f(sc, cam)
    N = fs["num_rendered"]
    cap = N + 1000
    fn = _nosync_forward(sc, cam, surfel, cap)
    torch.cuda.synchronize()
    isx = inter(fs)
    longest = int((isx["ranges"][:, 1] - isx["ranges"][:, 0]).max())
    assert fn["status"].tolist()[:3] == [N, longest, 0]
    if "mid" in name:
        assert 1024 < longest <= 8192
    if "long" in name:
        assert longest > 8192
    for k in (("color", "others", "radii") if surfel else ("color", "depth", "alpha", "radii")):
        assert torch.equal(fn[k], fs[k]), k
    inx = inter(fn)                                   # laid out for `cap` instances; the first N are the lists
    assert torch.equal(inx["ranges"], isx["ranges"])
    assert torch.equal(inx["point_list"][:N], isx["point_list"]) and torch.equal(inx["point_list_keys"][:N], isx["point_list_keys"])
    assert torch.equal(inx["n_contrib"], isx["n_contrib"])
    # backward on the no-sync buffers (R = capacity) against the backward on the stock buffers: atomics noise only
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 3, channels=(3, 7) if surfel else (3, 1, 1))]
    gs_runs = [bwd_f(sc, cam, fs, grads) for _ in range(4)]
    gn = bwd_f(sc, cam, fn, grads)
    for k in gn:
        assert h.close_to_runs(gn[k], [g[k] for g in gs_runs])[0], k


@pytest.mark.parametrize("surfel", [False, True], ids=["3dgs", "2dgs"])
def test_nosync_overflow_is_flagged_and_harmless(surfel):
    sc = h.make("C4" if surfel else "C1", P=30_000, W=200, H=136, random_bg=True); cam = sc.cameras[0]
    fs = (h.ours_surfel_forward if surfel else h.ours_forward)(sc, cam)
    N = fs["num_rendered"]
    fn = _nosync_forward(sc, cam, surfel, N - 1)     # one instance short
    torch.cuda.synchronize()
    st = fn["status"].tolist()
    assert st[0] == N and st[2] == 1
    # nothing was blended: the colour is the background, and a backward on these buffers writes zeros and faults nothing
    assert torch.equal(fn["color"], sc.bg.view(3, 1, 1).expand(3, cam.H, cam.W))
    grads = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 3, channels=(3, 7) if surfel else (3, 1, 1))]
    g = (h.ours_surfel_backward if surfel else h.ours_backward)(sc, cam, fn, grads)
    torch.cuda.synchronize()
    assert float(g["means3D"].abs().max()) == 0.0 and float(g["sh"].abs().max()) == 0.0
    # exactly enough capacity is fine
    fe = _nosync_forward(sc, cam, surfel, N)
    torch.cuda.synchronize()
    assert fe["status"].tolist()[2] == 0 and torch.equal(fe["color"], fs["color"])


def _public(mod, sc, cam, params, m2d):
    rs = mod.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, sc.sh_degree, cam.campos, False, False)
    return mod.GaussianRasterizer(rs)(params[0], m2d, params[2], shs=params[1], scales=params[3], rotations=params[4])


@pytest.mark.parametrize("surfel", [False, True], ids=["3dgs", "2dgs"])
def test_public_api_under_capacity_enable(surfel):
    if surfel:
        import scorp_b200.diff_surfel_rasterization as mod
    else:
        import scorp_b200.diff_gaussian_rasterization as mod
    sc = h.make("C4" if surfel else "C2", P=40_000, W=320, H=208, n_views=3)
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    plain = [_public(mod, sc, cam, params, torch.zeros_like(params[0])) for cam in sc.cameras]
    capacity.reset(); capacity.enable(margin=1.5)
    try:
        outs = []
        for cam in sc.cameras:                        # the first call learns the count (stock path), the others do not wait
            outs.append(_public(mod, sc, cam, params, torch.zeros_like(params[0], requires_grad=True)))
        torch.cuda.synchronize()
        assert not capacity.overflowed()
        for a, b in zip(outs, plain):
            for x, y in zip(a, b):
                assert torch.equal(x, y)
        # gradients through the no-sync forward
        imgs = [o for o in outs[1] if o.dtype.is_floating_point]
        torch.autograd.backward(imgs, [torch.ones_like(o) / o.numel() for o in imgs])
        g_nosync = [p.grad.clone() for p in params]
        for p in params:
            p.grad = None
        capacity.disable()
        eager = []                      # the stock path, four times: float atomics make every run a little different
        for _ in range(4):
            for p in params:
                p.grad = None
            imgs = [o for o in _public(mod, sc, sc.cameras[1], params, torch.zeros_like(params[0], requires_grad=True))
                    if o.dtype.is_floating_point]
            torch.autograd.backward(imgs, [torch.ones_like(o) / o.numel() for o in imgs])
            eager.append([p.grad.clone() for p in params])
        for i, a in enumerate(g_nosync):
            assert h.close_to_runs(a, [e[i] for e in eager])[0]
        # a view with more instances than the high-water mark allows: flagged, mark raised, the retry succeeds
        capacity.enable(margin=1.0, slack=0)
        key = ("2dgs" if surfel else "3dgs", sc.P, sc.cameras[0].W, sc.cameras[0].H)
        capacity._S.high[key] = 1000
        bad = _public(mod, sc, sc.cameras[0], params, torch.zeros_like(params[0]))
        torch.cuda.synchronize()
        assert capacity.overflowed()
        assert not torch.equal(bad[0], plain[0][0])
        with torch.no_grad():
            good = _public(mod, sc, sc.cameras[0], params, torch.zeros_like(params[0]))
        torch.cuda.synchronize()
        assert not capacity.overflowed() and torch.equal(good[0], plain[0][0])
        capacity._S.high[key] = 1000
        _public(mod, sc, sc.cameras[0], params, torch.zeros_like(params[0]))
        with pytest.raises(capacity.InstanceCapacityExceeded):
            capacity.check()
    finally:
        capacity.disable(); capacity.reset()


def test_forward_and_backward_captured_in_a_cuda_graph():
    """With the round trip gone a whole rasterizer step is capturable: replaying the graph after changing the camera and
    the parameters IN PLACE gives the images and gradients of an eager step on the new values."""
    import scorp_b200.diff_gaussian_rasterization as mod
    sc = h.make("C2", P=60_000, W=400, H=260, n_views=2)
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    cam_buf = torch.cat([sc.cameras[0].viewmatrix.flatten(), sc.cameras[0].projmatrix.flatten(), sc.cameras[0].campos]).clone()
    cam0 = sc.cameras[0]
    ups = [g.cuda() for g in synth.upstream_grads(cam0.W, cam0.H, 8)]

    def step():
        rs = mod.GaussianRasterizationSettings(cam0.H, cam0.W, cam0.tanfovx, cam0.tanfovy, sc.bg, 1.0, cam_buf[0:16].view(4, 4),
                                               cam_buf[16:32].view(4, 4), sc.sh_degree, cam_buf[32:35], False, False)
        outs = mod.GaussianRasterizer(rs)(params[0], torch.zeros_like(params[0], requires_grad=True), params[2], shs=params[1],
                                          scales=params[3], rotations=params[4])
        torch.autograd.backward([outs[0], outs[2], outs[3]], ups)
        return outs

    capacity.reset(); capacity.enable(margin=1.5)
    try:
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):                 # warm-up: learns the instance count, sets kernel attributes
            for _ in range(2):
                for p in params:
                    p.grad = None
                step()
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        assert not capacity.overflowed()
        for p in params:
            p.grad = None
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            g_outs = step()
        # new camera and perturbed parameters, in place; then replay
        cam1 = sc.cameras[1]
        with torch.no_grad():
            cam_buf.copy_(torch.cat([cam1.viewmatrix.flatten(), cam1.projmatrix.flatten(), cam1.campos]))
            params[0].add_(0.01); params[2].mul_(0.9)
        graph.replay()
        torch.cuda.synchronize()
        assert not capacity.overflowed()
        got = [o.clone() for o in g_outs] + [p.grad.clone() for p in params]
        # eager, stock path, same values
        capacity.disable()
        eager = []
        for _ in range(4):
            for p in params:
                p.grad = None
            e_outs = step()
            torch.cuda.synchronize()
            eager.append([p.grad.clone() for p in params])
        for a, b in zip(got[:4], e_outs):
            assert torch.equal(a, b)
        for i, a in enumerate(got[4:]):
            assert h.close_to_runs(a, [e[i] for e in eager])[0]
        assert float(got[4].abs().max()) > 0
    finally:
        capacity.disable(); capacity.reset(); capacity.forget_captured()
