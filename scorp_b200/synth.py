"""Seeded synthetic scenes of the shapes BASELINE.json names.

Tensors are generated on the CPU in fp32 from a seeded torch.Generator, so the product kernels,
the reference CUDA build and the CPU oracle all see identical bits. Camera matrices follow the
reference's conventions: world_view_transform / full_proj_transform are the TRANSPOSED 4x4s of
gs3dgs/scene/cameras.py:82-97 built like gs3dgs/utils/graphics_utils.py:38-71 (znear 0.01, zfar 100).
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch

# id -> (P, sh_degree, W, H, s0, views, mode)
CONFIGS = {
    "C1": dict(P=50_000, sh_degree=0, W=512, H=512, s0=0.03, seed=1001, surfel=False,
               desc="post_refine_gs object refinement: 50k Gaussians, SH0, 512x512, 1 view fwd+bwd"),
    "C2": dict(P=1_000_000, sh_degree=3, W=1297, H=840, s0=0.012, seed=1002, surfel=False,
               desc="train_3dgs scene: 1M Gaussians, SH3, 1297x840, fwd+bwd"),
    "C3": dict(P=6_000_000, sh_degree=3, W=1920, H=1080, s0=0.012, seed=1003, surfel=False,
               desc="large scene: 6M Gaussians, SH3, 1920x1080, 8 views sharded by camera"),
    "C4": dict(P=1_000_000, sh_degree=3, W=1297, H=840, s0=0.012, seed=1004, surfel=True,
               desc="train_2dgs: 1M surfels, SH3, 1297x840, colour + 7-channel map fwd+bwd"),
    "C5": dict(P=200_000, sh_degree=0, W=512, H=512, s0=0.03, seed=1005, surfel=False, forward_only=True, sweep_views=128,
               desc="align_3dgs pose sweep: 200k Gaussians, SH0, 512x512, forward only x128 rotations"),
}

CENTER = (0.0, 0.0, 6.0)


@dataclass
class Camera:
    W: int
    H: int
    tanfovx: float
    tanfovy: float
    viewmatrix: torch.Tensor      # (4,4) transposed world->view
    projmatrix: torch.Tensor      # (4,4) transposed full projection
    campos: torch.Tensor          # (3,)


@dataclass
class Scene:
    name: str
    P: int
    sh_degree: int
    means3D: torch.Tensor
    scales: torch.Tensor          # (P,3) 3DGS, (P,2) surfels; activated
    rotations: torch.Tensor       # (P,4) normalised (r,x,y,z)
    opacities: torch.Tensor       # (P,1) activated
    shs: torch.Tensor             # (P,M,3)
    bg: torch.Tensor              # (3,)
    cameras: list = field(default_factory=list)
    surfel: bool = False

    def to(self, device):
        for k in ("means3D", "scales", "rotations", "opacities", "shs", "bg"):
            setattr(self, k, getattr(self, k).to(device))
        for c in self.cameras:
            c.viewmatrix = c.viewmatrix.to(device); c.projmatrix = c.projmatrix.to(device)
            c.campos = c.campos.to(device)
        return self


def projection_matrix(znear, zfar, tanx, tany):
    """graphics_utils.py:51-71 restated with tan(fov/2) given directly."""
    P = torch.zeros(4, 4, dtype=torch.float32)
    right, top = tanx * znear, tany * znear
    P[0, 0] = 2.0 * znear / (2.0 * right)
    P[1, 1] = 2.0 * znear / (2.0 * top)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def make_camera(W, H, R=None, T=None, fx_factor=0.8):
    """R: camera-to-world rotation (as COLMAP readers store it), T: world-to-view translation."""
    R = np.eye(3) if R is None else np.asarray(R, dtype=np.float64)
    T = np.zeros(3) if T is None else np.asarray(T, dtype=np.float64)
    Rt = np.zeros((4, 4)); Rt[:3, :3] = R.T; Rt[:3, 3] = T; Rt[3, 3] = 1.0
    w2v = torch.tensor(np.float32(Rt))
    fx = fx_factor * W
    tanx = W / (2.0 * fx)
    tany = H / (2.0 * fx)
    view_t = w2v.transpose(0, 1).contiguous()
    proj_t = projection_matrix(0.01, 100.0, tanx, tany).transpose(0, 1)
    full = (view_t.unsqueeze(0).bmm(proj_t.unsqueeze(0))).squeeze(0).contiguous()
    campos = view_t.inverse()[3, :3].contiguous()
    return Camera(W, H, float(tanx), float(tany), view_t, full, campos)


def orbit_camera(W, H, Robj):
    """Camera seeing the scene as if the object were rotated by Robj about CENTER (kept at depth 6)."""
    Robj = np.asarray(Robj, dtype=np.float64)
    c = np.asarray(CENTER)
    # view = Robj (x - c) + c  => rotation part Robj, translation c - Robj c; R (c2w) = Robj^T
    return make_camera(W, H, R=Robj.T, T=c - Robj @ c)


def yaw(deg):
    a = math.radians(deg)
    return np.array([[math.cos(a), 0, math.sin(a)], [0, 1, 0], [-math.sin(a), 0, math.cos(a)]])


def random_rotations(n, seed=128):
    """n seeded uniform random rotation matrices (float64 (n,3,3)) -- a stand-in with the shape and role of the
    reference's rotation_matrices/rotations_128.npz, which does not travel to the GPU box."""
    g = torch.Generator(device="cpu").manual_seed(seed)
    q = torch.randn(n, 4, generator=g, dtype=torch.float64)
    q = q / q.norm(dim=1, keepdim=True)
    r, x, y, z = q.unbind(1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], 1).view(n, 3, 3)
    return R.numpy()


def make_scene(name="C1", P=None, n_views=1, W=None, H=None, sh_degree=None, seed=None, random_bg=False,
               rotations_npz=None, cameras=None):
    """Seeded scene. Generated with ONE intra-op thread: torch's vectorised CPU kernels (exp, sigmoid, norm) take a
    SIMD path for the body of a thread's chunk and a scalar path for its tail, and the two may differ in the last bit,
    so with N threads the values at the N-1 chunk boundaries depend on N. One thread makes the bits a function of the
    seed alone, on any host; bench.py prints a hash of the device-resident inputs (config.inputs_sha) to prove it."""
    nthreads = torch.get_num_threads()
    torch.set_num_threads(1)
    try:
        return _make_scene(name, P, n_views, W, H, sh_degree, seed, random_bg, rotations_npz, cameras)
    finally:
        torch.set_num_threads(nthreads)


def stress_camera(W, H, k):
    """Camera INSIDE the point cloud (at CENTER, looking along a direction that depends on k): about half of the
    Gaussians are behind it, many sit near the z = 0.2 cull plane with radii of tens of pixels, and most of the rest
    lie outside the frustum, i.e. on the +-1.3*tanfov clamp of computeCov2D (forward.cu:82-87)."""
    a, b = math.radians(37.0 * k), math.radians(11.0 * k)
    pitch = np.array([[1, 0, 0], [0, math.cos(b), -math.sin(b)], [0, math.sin(b), math.cos(b)]])
    Robj = pitch @ yaw(math.degrees(a))
    c = np.asarray(CENTER)
    # view = Robj (x - c): the camera sits at the cloud's centre
    return make_camera(W, H, R=Robj.T, T=-Robj @ c)


def _make_scene(name, P, n_views, W, H, sh_degree, seed, random_bg, rotations_npz, cameras):
    cfg = dict(CONFIGS[name])
    P = cfg["P"] if P is None else P
    W = cfg["W"] if W is None else W
    H = cfg["H"] if H is None else H
    D = cfg["sh_degree"] if sh_degree is None else sh_degree
    seed = cfg["seed"] if seed is None else seed
    g = torch.Generator(device="cpu").manual_seed(seed)
    M = (D + 1) ** 2
    c = torch.tensor(CENTER)
    means = c + (torch.rand(P, 3, generator=g) * 8.0 - 4.0)
    ns = 2 if cfg["surfel"] else 3
    scales = torch.exp(math.log(cfg["s0"]) + 0.6 * torch.randn(P, ns, generator=g))
    rot = torch.randn(P, 4, generator=g)
    rot = rot / rot.norm(dim=1, keepdim=True)
    opac = torch.sigmoid(2.0 * torch.randn(P, 1, generator=g))
    shs = torch.empty(P, M, 3)
    shs[:, 0, :] = (0.5 + 0.5 * torch.randn(P, 3, generator=g)) / 0.2821
    if M > 1:
        shs[:, 1:, :] = 0.1 * torch.randn(P, M - 1, 3, generator=g)
    bg = torch.rand(3, generator=g) if random_bg else torch.zeros(3)
    cams = list(cameras) if cameras is not None else []
    for v in range(n_views if cameras is None else 0):
        if v == 0:
            cams.append(make_camera(W, H))
        elif rotations_npz is not None:
            cams.append(orbit_camera(W, H, rotations_npz[v % len(rotations_npz)]))
        else:
            cams.append(orbit_camera(W, H, yaw(45.0 * v)))
    return Scene(name, P, D, means.contiguous(), scales.contiguous(), rot.contiguous(), opac.contiguous(),
                 shs.contiguous(), bg, cams, cfg["surfel"])


def upstream_grads(W, H, seed, channels=(3, 1, 1)):
    """Fixed upstream gradients N(0,1)/(W*H), one tensor per output."""
    g = torch.Generator(device="cpu").manual_seed(seed + 7919)
    return [torch.randn(ch, H, W, generator=g) / float(W * H) for ch in channels]
