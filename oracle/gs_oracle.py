"""TEST INFRASTRUCTURE (checker): ctypes/numpy driver of oracle/libgs_oracle.so, the plain-C CPU
restatement of the reference rasterizers (oracle/gs_oracle.c). Only tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / reference legs may import this; nothing under scorp_b200/ does."""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB = HERE / "libgs_oracle.so"
_lib = None


def build():
    subprocess.run(["make", "-s", "-C", str(HERE), "oracle"], check=True)


def lib():
    global _lib
    if _lib is None:
        if not LIB.exists():
            build()
        _lib = C.CDLL(str(LIB))
        _lib.gs_oracle_forward.restype = C.c_void_p
        _lib.gs_oracle_num_rendered.restype = C.c_int64
        _lib.gs_oracle_num_rendered.argtypes = [C.c_void_p]
        _lib.gs_oracle_array.restype = C.c_void_p
        _lib.gs_oracle_array.argtypes = [C.c_void_p, C.c_int]
        _lib.gs_oracle_free.argtypes = [C.c_void_p]
    return _lib


def _f(a):
    """contiguous float32 numpy (or None for absent/empty)."""
    if a is None:
        return None
    if hasattr(a, "detach"):
        a = a.detach().cpu().numpy()
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a if a.size else None


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


_ARRAYS = dict(depths=(0, np.float32, 1), means2D=(1, np.float32, 2), cov3D=(2, np.float32, 6),
               conic_opacity=(3, np.float32, 4), rgb=(4, np.float32, 3), clamped=(5, np.uint8, 3),
               tiles_touched=(6, np.uint32, 1), point_offsets=(7, np.uint32, 1))


class Forward:
    """Result of the oracle forward: outputs, intermediates and the state needed by backward."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def __del__(self):
        st = self.__dict__.get("state")
        if st and _lib is not None:
            try:
                _lib.gs_oracle_free(C.c_void_p(st))
            except Exception:
                pass
            self.state = None


def forward(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, cov3D_precomp, view, proj,
            tanfovx, tanfovy, H, W, shs, degree, campos, threads=None):
    L = lib()
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    bg, means3D, colors_precomp, opacities = _f(bg), _f(means3D), _f(colors_precomp), _f(opacities)
    scales, rotations, cov3D_precomp, view, proj, campos, shs = map(_f, (scales, rotations, cov3D_precomp, view, proj, campos, shs))
    P = 0 if means3D is None else means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    color = np.zeros((3, H, W), np.float32); depth = np.zeros((1, H, W), np.float32)
    alpha = np.zeros((1, H, W), np.float32); radii = np.zeros((P,), np.int32)
    if P == 0:      # the torch glue skips the rasterizer entirely: zero images (rasterize_points.cu:83)
        return Forward(state=None, color=color, depth=depth, alpha=alpha, radii=radii, num_rendered=0, inter={},
                       P=0, M=M, H=H, W=W, degree=degree, scale_modifier=scale_modifier, tanfovx=tanfovx,
                       tanfovy=tanfovy, keep={})
    st = L.gs_oracle_forward(C.c_int(P), C.c_int(degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H), _p(means3D),
                             _p(shs), _p(colors_precomp), _p(opacities), _p(scales), C.c_float(scale_modifier),
                             _p(rotations), _p(cov3D_precomp), _p(view), _p(proj), _p(campos), C.c_float(tanfovx),
                             C.c_float(tanfovy), _p(color), _p(depth), _p(alpha), _p(radii))
    if not st:
        raise RuntimeError("oracle forward failed (need shs or colors_precomp)")
    N = int(L.gs_oracle_num_rendered(st))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def arr(which, dtype, shape):
        n = int(np.prod(shape))
        if n == 0:
            return np.zeros(shape, dtype)
        ptr = L.gs_oracle_array(st, which)
        buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()

    inter = {k: arr(w, dt, (P, c) if c > 1 else (P,)) for k, (w, dt, c) in _ARRAYS.items()}
    inter["point_list_keys"] = arr(8, np.uint64, (N,))
    inter["point_list"] = arr(9, np.uint32, (N,))
    inter["ranges"] = arr(10, np.uint32, (tiles, 2))
    inter["n_contrib"] = arr(11, np.uint32, (H, W))
    keep = dict(bg=bg, means3D=means3D, colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                cov3D_precomp=cov3D_precomp, view=view, proj=proj, campos=campos, shs=shs)
    return Forward(state=st, color=color, depth=depth, alpha=alpha, radii=radii, num_rendered=N, inter=inter,
                   P=P, M=M, H=H, W=W, degree=degree, scale_modifier=scale_modifier, tanfovx=tanfovx,
                   tanfovy=tanfovy, keep=keep)


def backward(fwd: Forward, dL_dcolor, dL_ddepth, dL_dalpha):
    L = lib()
    k = fwd.keep
    P, M = fwd.P, fwd.M
    g = dict(means2D=np.zeros((P, 3), np.float32), conic=np.zeros((P, 4), np.float32),
             opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
             depths=np.zeros((P, 1), np.float32), means3D=np.zeros((P, 3), np.float32),
             cov3D=np.zeros((P, 6), np.float32), sh=np.zeros((P, M, 3), np.float32),
             scales=np.zeros((P, 3), np.float32), rotations=np.zeros((P, 4), np.float32))
    dc, dd, da = _f(dL_dcolor), _f(dL_ddepth), _f(dL_dalpha)
    L.gs_oracle_backward(C.c_void_p(fwd.state), C.c_int(fwd.degree), C.c_int(M), _p(k["bg"]), _p(k["means3D"]),
                         _p(k["shs"]), _p(k["colors_precomp"]), _p(fwd.alpha), _p(k["scales"]),
                         C.c_float(fwd.scale_modifier), _p(k["rotations"]), _p(k["cov3D_precomp"]), _p(k["view"]),
                         _p(k["proj"]), _p(k["campos"]), C.c_float(fwd.tanfovx), C.c_float(fwd.tanfovy),
                         _p(dc), _p(dd), _p(da), _p(g["means2D"]), _p(g["conic"]), _p(g["opacity"]), _p(g["colors"]),
                         _p(g["depths"]), _p(g["means3D"]), _p(g["cov3D"]), _p(g["sh"]) if M else None,
                         _p(g["scales"]), _p(g["rotations"]))
    return g


def mark_visible(means3D, view, proj):
    L = lib()
    m, v, p = _f(means3D), _f(view), _f(proj)
    out = np.zeros((m.shape[0],), np.uint8)
    L.gs_oracle_mark_visible(C.c_int(m.shape[0]), _p(m), _p(v), _p(p), _p(out))
    return out.astype(bool)


# ---- 2DGS surfels (oracle/surfel_oracle.c) ---------------------------------------------------------

_S_ARRAYS = dict(depths=(0, np.float32, 1), means2D=(1, np.float32, 2), transMat=(2, np.float32, 9),
                 normal_opacity=(3, np.float32, 4), rgb=(4, np.float32, 3), clamped=(5, np.uint8, 3),
                 tiles_touched=(6, np.uint32, 1), point_offsets=(7, np.uint32, 1))


class SurfelForward(Forward):
    def __del__(self):
        st = self.__dict__.get("state")
        if st and _lib is not None:
            try:
                _lib.surfel_oracle_free(C.c_void_p(st))
            except Exception:
                pass
            self.state = None


def _slib():
    L = lib()
    if not getattr(L, "_surfel_bound", False):
        L.surfel_oracle_forward.restype = C.c_void_p
        L.surfel_oracle_num_rendered.restype = C.c_int64
        L.surfel_oracle_num_rendered.argtypes = [C.c_void_p]
        L.surfel_oracle_array.restype = C.c_void_p
        L.surfel_oracle_array.argtypes = [C.c_void_p, C.c_int]
        L.surfel_oracle_free.argtypes = [C.c_void_p]
        L._surfel_bound = True
    return L


def surfel_forward(bg, means3D, colors_precomp, opacities, scales, rotations, scale_modifier, transMat_precomp,
                   view, proj, tanfovx, tanfovy, H, W, shs, degree, campos, threads=None):
    """Restatement of the surfel Rasterizer::forward: color[3,H,W], others[7,H,W], radii + every intermediate."""
    L = _slib()
    if threads:
        os.environ["OMP_NUM_THREADS"] = str(threads)
    bg, means3D, colors_precomp, opacities = _f(bg), _f(means3D), _f(colors_precomp), _f(opacities)
    scales, rotations, transMat_precomp, view, proj, campos, shs = map(
        _f, (scales, rotations, transMat_precomp, view, proj, campos, shs))
    P = 0 if means3D is None else means3D.shape[0]
    M = 0 if shs is None else shs.shape[1]
    color = np.zeros((3, H, W), np.float32); others = np.zeros((7, H, W), np.float32)
    radii = np.zeros((P,), np.int32)
    base = dict(P=P, M=M, H=H, W=W, degree=degree, scale_modifier=scale_modifier, tanfovx=tanfovx, tanfovy=tanfovy)
    if P == 0:
        return SurfelForward(state=None, color=color, others=others, radii=radii, num_rendered=0, inter={}, keep={}, **base)
    st = L.surfel_oracle_forward(C.c_int(P), C.c_int(degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H), _p(means3D),
                                 _p(shs), _p(colors_precomp), _p(opacities), _p(scales), C.c_float(scale_modifier),
                                 _p(rotations), _p(transMat_precomp), _p(view), _p(proj), _p(campos),
                                 C.c_float(tanfovx), C.c_float(tanfovy), _p(color), _p(others), _p(radii))
    if not st:
        raise RuntimeError("surfel oracle forward failed (need shs or colors_precomp)")
    N = int(L.surfel_oracle_num_rendered(st))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)

    def arr(which, dtype, shape):
        n = int(np.prod(shape))
        if n == 0:
            return np.zeros(shape, dtype)
        ptr = L.surfel_oracle_array(st, which)
        buf = (C.c_char * (n * np.dtype(dtype).itemsize)).from_address(ptr)
        return np.frombuffer(buf, dtype=dtype).reshape(shape).copy()

    inter = {k: arr(w, dt, (P, c) if c > 1 else (P,)) for k, (w, dt, c) in _S_ARRAYS.items()}
    inter["point_list_keys"] = arr(8, np.uint64, (N,))
    inter["point_list"] = arr(9, np.uint32, (N,))
    inter["ranges"] = arr(10, np.uint32, (tiles, 2))
    inter["n_contrib"] = arr(11, np.uint32, (H, W)); inter["median_contrib"] = arr(12, np.uint32, (H, W))
    inter["final_T"] = arr(13, np.float32, (H, W)); inter["M1"] = arr(14, np.float32, (H, W))
    inter["M2"] = arr(15, np.float32, (H, W))
    keep = dict(bg=bg, means3D=means3D, colors_precomp=colors_precomp, scales=scales, rotations=rotations,
                transMat_precomp=transMat_precomp, view=view, proj=proj, campos=campos, shs=shs)
    return SurfelForward(state=st, color=color, others=others, radii=radii, num_rendered=N, inter=inter, keep=keep, **base)


def surfel_backward(fwd: SurfelForward, dL_dcolor, dL_dothers):
    L = _slib()
    k = fwd.keep
    P, M = fwd.P, fwd.M
    g = dict(means2D=np.zeros((P, 3), np.float32), normal=np.zeros((P, 3), np.float32),
             opacity=np.zeros((P, 1), np.float32), colors=np.zeros((P, 3), np.float32),
             means3D=np.zeros((P, 3), np.float32), transMat=np.zeros((P, 9), np.float32),
             sh=np.zeros((P, M, 3), np.float32), scales=np.zeros((P, 2), np.float32),
             rotations=np.zeros((P, 4), np.float32))
    if P == 0:
        return g
    dc, do = _f(dL_dcolor), _f(dL_dothers)
    L.surfel_oracle_backward(C.c_void_p(fwd.state), C.c_int(fwd.degree), C.c_int(M), _p(k["bg"]), _p(k["means3D"]),
                             _p(k["shs"]), _p(k["colors_precomp"]), _p(k["scales"]), C.c_float(fwd.scale_modifier),
                             _p(k["rotations"]), _p(k["transMat_precomp"]), _p(k["view"]), _p(k["proj"]),
                             _p(k["campos"]), C.c_float(fwd.tanfovx), C.c_float(fwd.tanfovy), _p(dc), _p(do),
                             _p(g["means2D"]), _p(g["normal"]), _p(g["opacity"]), _p(g["colors"]), _p(g["means3D"]),
                             _p(g["transMat"]), _p(g["sh"]) if M else None, _p(g["scales"]), _p(g["rotations"]))
    return g


# ---- simple-knn (oracle/knn_oracle.c) ---------------------------------------------------------------

def knn_dist2(points, brute=False):
    """Restatement of SimpleKNN::knn (Morton order + 1024-point boxes), or the exhaustive cross-check."""
    L = lib()
    pts = _f(points)
    P = 0 if pts is None else pts.shape[0]
    out = np.zeros((P,), np.float32)
    if P:
        (L.knn_oracle_brute if brute else L.knn_oracle)(C.c_int(P), _p(pts), _p(out))
    return out


# ---- photometric loss (oracle/loss_oracle.c) --------------------------------------------------------

def l1_ssim(img, gt, lambda_dssim=0.2, grads=True):
    """Restatement of loss_utils.py l1_loss / ssim and train_3dgs.py:107.
    -> dict(loss, l1, ssim [, dl1, dssim, dloss]) with the gradients w.r.t. img."""
    L = lib()
    a, b = _f(img), _f(gt)
    Cn, H, W = a.shape
    out = (C.c_double * 3)()
    dl1 = np.zeros_like(a) if grads else None
    dss = np.zeros_like(a) if grads else None
    L.loss_oracle_l1_ssim(C.c_int(Cn), C.c_int(H), C.c_int(W), _p(a), _p(b), C.c_float(lambda_dssim), out, _p(dl1), _p(dss))
    res = dict(loss=out[0], l1=out[1], ssim=out[2])
    if grads:
        res.update(dl1=dl1, dssim=dss, dloss=((1.0 - lambda_dssim) * dl1 - lambda_dssim * dss).astype(np.float32))
    return res
