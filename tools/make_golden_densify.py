"""Golden vectors for densify_and_prune: the reference's own GaussianModel
(/root/reference/gs3dgs/scene/gaussian_model.py and gs2dgs/scene/gaussian_model.py, imported unmodified) run on CPU
fp32 in this container. The module's imports that are absent here (plyfile, simple_knn._C) are stubbed -- densification
does not touch them -- and the `device="cuda"` literals of its torch.zeros / torch.ones calls are dropped by a wrapper
for the duration of the run. The normal samples the reference draws are recovered by replaying the generator state with
the same shape (torch.normal(mean, std) = normal_(0, 1) * std + mean); tests/test_oracle_densify_golden.py confirms that
they reproduce the reference's children.
Writes tests/golden/densify.npz.   python tools/make_golden_densify.py"""
import importlib
import sys
import types
from pathlib import Path

import numpy as np
import torch

sys.dont_write_bytecode = True
ROOT = Path(__file__).resolve().parent.parent
NAMES = (("xyz", "_xyz"), ("f_dc", "_features_dc"), ("f_rest", "_features_rest"), ("opacity", "_opacity"),
         ("scaling", "_scaling"), ("rotation", "_rotation"))


def _reference(pkg):
    sys.path.insert(0, "/root/reference")
    for name in ("plyfile", "simple_knn", "simple_knn._C"):
        sys.modules.setdefault(name, types.ModuleType(name))
    sys.modules["plyfile"].PlyData = sys.modules["plyfile"].PlyElement = object
    sys.modules["simple_knn._C"].distCUDA2 = None
    return importlib.import_module(f"{pkg}.scene.gaussian_model").GaussianModel


class _cpu_literals:
    """drop device="cuda" from the factory calls while the reference code runs"""
    FNS = ("zeros", "ones", "empty", "tensor", "zeros_like", "ones_like")

    def __enter__(self):
        self.saved = {f: getattr(torch, f) for f in self.FNS}
        for f, fn in self.saved.items():
            setattr(torch, f, (lambda fn: lambda *a, **k: fn(*a, **{x: y for x, y in k.items() if x != "device"}))(fn))
        self.cache = torch.cuda.empty_cache
        torch.cuda.empty_cache = lambda: None

    def __exit__(self, *exc):
        for f, fn in self.saved.items():
            setattr(torch, f, fn)
        torch.cuda.empty_cache = self.cache


def case(pkg, P, sh_rest, scale_dim, seed, max_grad, min_opacity, extent, max_screen_size, percent_dense, with_state=True):
    GaussianModel = _reference(pkg)
    g = torch.Generator().manual_seed(seed)
    pc = GaussianModel(3)
    pc._xyz = torch.nn.Parameter(torch.randn(P, 3, generator=g) * 2)
    pc._features_dc = torch.nn.Parameter(torch.randn(P, 1, 3, generator=g))
    pc._features_rest = torch.nn.Parameter(torch.randn(P, sh_rest, 3, generator=g) * 0.1)
    pc._opacity = torch.nn.Parameter(torch.randn(P, 1, generator=g) * 3)               # sigmoid spans the 0.005 cull
    pc._scaling = torch.nn.Parameter(torch.randn(P, scale_dim, generator=g) * 1.2 - 3.0)  # exp spans both extents
    pc._rotation = torch.nn.Parameter(torch.randn(P, 4, generator=g) * 1.5)            # un-normalised, as stored
    with _cpu_literals():
        pc.training_setup(types.SimpleNamespace(percent_dense=percent_dense, position_lr_init=0.00016, position_lr_final=1.6e-6,
                                                position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025,
                                                opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001))
    pc.max_radii2D = torch.randint(0, 60, (P,), generator=g).float()
    denom = torch.randint(0, 4, (P, 1), generator=g).float()                       # zeros -> 0/0 -> NaN -> 0
    pc.denom = denom.clone()
    pc.xyz_gradient_accum = torch.rand(P, 1, generator=g) * denom * (2.0 * max_grad)
    if with_state:          # one optimizer step so the moments are non-trivial
        for _, attr in NAMES:
            p = getattr(pc, attr)
            p.grad = torch.randn(p.shape, generator=g) * 1e-3
        pc.optimizer.step()
        pc.optimizer.zero_grad(set_to_none=True)
    out = {}
    for n, attr in NAMES:
        out[f"in_{n}"] = getattr(pc, attr).detach().numpy().copy()
        st = pc.optimizer.state.get(getattr(pc, attr), None)
        if st:
            out[f"in_m_{n}"] = st["exp_avg"].numpy().copy(); out[f"in_v_{n}"] = st["exp_avg_sq"].numpy().copy()
    out["in_accum"] = pc.xyz_gradient_accum.numpy().copy(); out["in_denom"] = pc.denom.numpy().copy()

    torch.manual_seed(seed + 7)
    rng = torch.get_rng_state()
    drawn = []
    real_normal = torch.normal

    def spy(mean, std, **kw):          # record the shape the reference asked for; the call itself is untouched
        drawn.append(tuple(std.shape))
        return real_normal(mean=mean, std=std, **kw)
    torch.normal = spy
    try:
        with _cpu_literals():
            pc.densify_and_prune(max_grad, min_opacity, extent, max_screen_size)
    finally:
        torch.normal = real_normal
    assert len(drawn) == 1
    torch.set_rng_state(rng)
    normals = torch.empty(drawn[0]).normal_()
    for n, attr in NAMES:
        out[f"out_{n}"] = getattr(pc, attr).detach().numpy().copy()
        st = pc.optimizer.state.get(getattr(pc, attr), None)
        if st:
            out[f"out_m_{n}"] = st["exp_avg"].numpy().copy(); out[f"out_v_{n}"] = st["exp_avg_sq"].numpy().copy()
    out["normals"] = normals.numpy()
    out["cfg"] = np.array([max_grad, min_opacity, extent, max_screen_size or 0.0, percent_dense], np.float64)
    out["out_stats_rows"] = np.int32([pc.xyz_gradient_accum.shape[0], pc.denom.shape[0], pc.max_radii2D.shape[0]])
    assert float(pc.xyz_gradient_accum.abs().max()) == 0 and float(pc.max_radii2D.abs().max()) == 0
    return out


def reset_case(P=900, seed=21):
    """GaussianModel.reset_opacity (gaussian_model.py:253-256) after one optimizer step"""
    GaussianModel = _reference("gs3dgs")
    g = torch.Generator().manual_seed(seed)
    pc = GaussianModel(3)
    pc._xyz = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    pc._features_dc = torch.nn.Parameter(torch.randn(P, 1, 3, generator=g))
    pc._features_rest = torch.nn.Parameter(torch.randn(P, 15, 3, generator=g))
    pc._opacity = torch.nn.Parameter(torch.randn(P, 1, generator=g) * 4)
    pc._scaling = torch.nn.Parameter(torch.randn(P, 3, generator=g))
    pc._rotation = torch.nn.Parameter(torch.randn(P, 4, generator=g))
    with _cpu_literals():
        pc.training_setup(types.SimpleNamespace(percent_dense=0.01, position_lr_init=0.00016, position_lr_final=1.6e-6,
                                                position_lr_delay_mult=0.01, position_lr_max_steps=30000, feature_lr=0.0025,
                                                opacity_lr=0.05, scaling_lr=0.005, rotation_lr=0.001))
    pc._opacity.grad = torch.randn(P, 1, generator=g) * 1e-3
    pc.optimizer.step()
    out = {"in_opacity": pc._opacity.detach().numpy().copy()}
    step = pc.optimizer.state[pc._opacity]["step"]
    with _cpu_literals():
        pc.reset_opacity()
    st = pc.optimizer.state[pc._opacity]
    assert st["step"] is step and float(st["exp_avg"].abs().max()) == 0 and float(st["exp_avg_sq"].abs().max()) == 0
    out["out_opacity"] = pc._opacity.detach().numpy().copy()
    return out


def main():
    cases = {
        "a": case("gs3dgs", 1200, 15, 3, 11, 0.0002, 0.005, 4.0, 20, 0.01),
        "b": case("gs3dgs", 800, 0, 3, 12, 0.0002, 0.005, 6.0, None, 0.01),        # no size prune; SH0 model
        "c": case("gs2dgs", 1000, 8, 2, 13, 0.0002, 0.05, 3.0, 20, 0.01),          # 2DGS: two scales, third std is 0
        "d": case("gs3dgs", 600, 3, 3, 14, 0.0002, 0.005, 5.0, 20, 0.01, with_state=False),   # no optimizer state yet
    }
    flat = {f"{c}_{k}": v for c, d in cases.items() for k, v in d.items()}
    flat.update({f"reset_{k}": v for k, v in reset_case().items()})
    np.savez_compressed(ROOT / "tests" / "golden" / "densify.npz", cases=np.array(sorted(cases)), **flat)
    for c, d in cases.items():
        print(c, "P", d["in_xyz"].shape[0], "->", d["out_xyz"].shape[0], "rows; samples", d["normals"].shape)
    print("wrote tests/golden/densify.npz")


if __name__ == "__main__":
    main()
