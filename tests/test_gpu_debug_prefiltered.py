"""`debug=True` and `prefiltered=True` of GaussianRasterizationSettings, both packages.

Reference behaviour: with `debug`, the Python layer deep-copies the arguments to the CPU before the call and writes
them to snapshot_fw.dump / snapshot_bw.dump if the call raises (DGR diff_gaussian_rasterization/__init__.py:83-90,
135-142; DSR the same), and the C++ layer synchronises and checks for errors after every stage
(cuda_rasterizer/auxiliary.h:166-173). `prefiltered` promises that no point is culled by the frustum test; a culled
point then executes __trap() (auxiliary.h:154-161).

A false `prefiltered` promise is not provoked on purpose (a device-side trap is an Xid fault on the GPU box); the
dump path is driven by failures that do not fault the device."""
import os
from pathlib import Path

import pytest
import torch

import helpers as h
from scorp_b200 import synth

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def _mods(surfel):
    if surfel:
        import scorp_b200.diff_surfel_rasterization as mod
    else:
        import scorp_b200.diff_gaussian_rasterization as mod
    return mod


def _render(mod, sc, cam, prefiltered, debug, backward=True):
    rs = mod.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, sc.sh_degree, cam.campos, prefiltered, debug)
    params = [t.clone().requires_grad_(True) for t in (sc.means3D, sc.shs, sc.opacities, sc.scales, sc.rotations)]
    m2d = torch.zeros_like(params[0], requires_grad=True)
    outs = mod.GaussianRasterizer(rs)(params[0], m2d, params[2], shs=params[1], scales=params[3], rotations=params[4])
    if backward:
        imgs = [o for o in outs if o.dtype.is_floating_point]
        gr = [g.cuda() for g in synth.upstream_grads(cam.W, cam.H, 4, channels=tuple(o.shape[0] for o in imgs))]
        torch.autograd.backward(imgs, gr)
    return outs, [p.grad for p in params] + [m2d.grad]


@pytest.mark.parametrize("surfel", [False, True], ids=["3dgs", "2dgs"])
def test_debug_and_prefiltered_do_not_change_results(surfel):
    """All Gaussians of camera 0 are in front of the near plane (view z in [2, 10]), so prefiltered=True is a true
    promise; debug=True takes the per-stage synchronisation path. Same images and radii as the plain call and as the
    reference CUDA; gradients equal up to the order of the float atomics."""
    mod = _mods(surfel)
    sc = h.make("C4" if surfel else "C2", P=30_000, W=300, H=200, random_bg=True); cam = sc.cameras[0]
    plain, g_plain = _render(mod, sc, cam, False, False)
    g_runs = [g_plain] + [_render(mod, sc, cam, False, False)[1] for _ in range(3)]      # run-to-run spread of the atomics
    for prefiltered, debug in ((True, False), (False, True), (True, True)):
        outs, grads = _render(mod, sc, cam, prefiltered, debug)
        for a, b in zip(outs, plain):
            assert torch.equal(a, b), (prefiltered, debug)
        for i, a in enumerate(grads):
            assert h.close_to_runs(a, [g[i] for g in g_runs])[0], (prefiltered, debug)
    from oracle import ref_cuda
    if ref_cuda.available("dsr" if surfel else "dgr"):
        fr = (h.ref_surfel_forward if surfel else h.ref_forward)(sc, cam)
        assert torch.equal(plain[0], fr["color"]) and torch.equal(plain[1], fr["radii"])
    assert not os.path.exists("snapshot_fw.dump") and not os.path.exists("snapshot_bw.dump")


@pytest.mark.parametrize("surfel", [False, True], ids=["3dgs", "2dgs"])
def test_debug_dumps_the_inputs_when_a_call_fails(surfel, tmp_path, monkeypatch):
    """Forced failures with debug=True: the forward is given an SH layout the library rejects (M = 2 coefficients), the
    backward is made to fail by a stubbed `_C` entry point; both must leave the reference's snapshot files holding the
    CPU copies of the argument tuple, print the reference's message and re-raise. (A false `prefiltered` promise would
    also fail the forward, through __trap(); that is deliberately NOT provoked here: a device-side trap is an Xid fault
    on the GPU box.)"""
    from scorp_b200 import _lib
    mod = _mods(surfel)
    monkeypatch.chdir(tmp_path)
    sc = h.make("C4" if surfel else "C1", P=3000, W=160, H=120); cam = sc.cameras[0]
    rs = mod.GaussianRasterizationSettings(cam.H, cam.W, cam.tanfovx, cam.tanfovy, sc.bg, 1.0, cam.viewmatrix,
                                           cam.projmatrix, 0, cam.campos, False, True)
    bad_sh = torch.zeros(sc.P, 2, 3, device="cuda")
    with pytest.raises(_lib.GsrError, match="SH"):
        mod.GaussianRasterizer(rs)(sc.means3D, torch.zeros_like(sc.means3D), sc.opacities, shs=bad_sh, scales=sc.scales,
                                   rotations=sc.rotations)
    saved = torch.load(tmp_path / "snapshot_fw.dump", weights_only=False)
    # the argument tuple of _C.rasterize_gaussians, tensors copied to the CPU before the call (reference :83-84)
    assert len(saved) == 19 and all((not isinstance(a, torch.Tensor)) or a.device.type == "cpu" for a in saved)
    assert saved[1].shape == (sc.P, 3) and saved[14].shape == (sc.P, 2, 3) and saved[18] is True
    assert torch.equal(saved[1], sc.means3D.cpu())

    # backward: a good forward, then the `_C` backward fails
    means = sc.means3D.clone().requires_grad_(True)
    outs = mod.GaussianRasterizer(rs)(means, torch.zeros_like(means), sc.opacities, shs=sc.shs, scales=sc.scales,
                                      rotations=sc.rotations)
    def boom(*a, **k):
        raise RuntimeError("forced backward failure")
    monkeypatch.setattr(mod._C, "rasterize_gaussians_backward", boom)
    with pytest.raises(RuntimeError, match="forced backward failure"):
        outs[0].sum().backward()
    saved_b = torch.load(tmp_path / "snapshot_bw.dump", weights_only=False)
    assert all((not isinstance(a, torch.Tensor)) or a.device.type == "cpu" for a in saved_b)
    assert any(isinstance(a, torch.Tensor) and a.shape == (3, cam.H, cam.W) for a in saved_b)      # the upstream colour gradient
