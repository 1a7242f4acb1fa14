"""TEST INFRASTRUCTURE (checker), not product code.

ctypes driver for oracle/_ref/libref_{dgr,dsr,knn}.so: the UNMODIFIED reference CUDA sources
compiled for sm_100a by oracle/Makefile behind small C-ABI shims (oracle/ref_shim_*.cu).
Used by tests/ (golden vectors, GPU parity), __graft_entry__.smoke() and bench.py's reference
arm. Nothing under scorp_b200/ imports this.

The reference launches on the legacy default stream; callers synchronise around these calls.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

REF_DIR = Path(__file__).resolve().parent / "_ref"
_ALLOC = C.CFUNCTYPE(C.c_void_p, C.c_void_p, C.c_int, C.c_size_t)
_libs = {}


def available(which="dgr") -> bool:
    return (REF_DIR / f"libref_{which}.so").exists()


def _load(which):
    if which not in _libs:
        _libs[which] = C.CDLL(str(REF_DIR / f"libref_{which}.so"))
    return _libs[which]


def _p(t):
    return None if t is None or t.numel() == 0 else C.c_void_p(t.data_ptr())


class _Arena:
    def __init__(self, device):
        self.device, self.bufs = device, {}
        self.fn = _ALLOC(self._alloc)

    def _alloc(self, _ctx, which, n):
        t = torch.empty(max(int(n), 1) + 256, dtype=torch.uint8, device=self.device)
        self.bufs[which] = t
        return t.data_ptr()


def _layout(lib, name, n, count):
    off = (C.c_size_t * count)()
    total = C.c_size_t(0)
    getattr(lib, name)(C.c_size_t(n), off, C.byref(total))
    return list(off), total.value


def _slice(buf, base_off, nbytes, dtype, shape):
    """view of a scratch sub-array; the reference aligns sub-arrays to 128 B from the buffer base."""
    base = buf.data_ptr()
    a = (128 - base % 128) % 128 if base % 128 else 0
    raw = buf[a + base_off: a + base_off + nbytes]
    return raw.view(dtype).view(shape)


def dgr_forward(bg, means3D, colors, opacity, scales, rotations, scale_modifier, cov3D, view, proj,
                tanfovx, tanfovy, H, W, sh, degree, campos, prefiltered=False, debug=False, sync=True):
    """Reference Rasterizer::forward. Returns dict with outputs, buffers and sliced intermediates."""
    lib = _load("dgr")
    lib.ref_dgr_forward.restype = C.c_int
    dev = means3D.device
    P = means3D.shape[0]
    M = sh.shape[1] if sh is not None and sh.numel() else 0
    color = torch.zeros(3, H, W, device=dev); depth = torch.zeros(1, H, W, device=dev)
    alpha = torch.zeros(1, H, W, device=dev); radii = torch.zeros(P, dtype=torch.int32, device=dev)
    ar = _Arena(dev)
    if sync:
        torch.cuda.synchronize(dev)
    n = lib.ref_dgr_forward(ar.fn, None, C.c_int(P), C.c_int(degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H),
                            _p(means3D), _p(sh), _p(colors), _p(opacity), _p(scales), C.c_float(scale_modifier),
                            _p(rotations), _p(cov3D), _p(view), _p(proj), _p(campos), C.c_float(tanfovx),
                            C.c_float(tanfovy), C.c_int(int(prefiltered)), _p(color), _p(depth), _p(alpha),
                            _p(radii), C.c_int(int(debug)))
    if sync:
        torch.cuda.synchronize(dev)
    if n < 0:
        raise RuntimeError("reference forward raised")
    return dict(num_rendered=n, color=color, depth=depth, alpha=alpha, radii=radii,
                geom=ar.bufs.get(0), binning=ar.bufs.get(1), image=ar.bufs.get(2), P=P, M=M, H=H, W=W)


def dgr_intermediates(out):
    """Slice the reference's scratch buffers (rasterizer_impl.h:23-73) into named tensors."""
    lib = _load("dgr")
    P, N, H, W = out["P"], out["num_rendered"], out["H"], out["W"]
    g_off, _ = _layout(lib, "ref_dgr_geom_layout", P, 10)
    i_off, _ = _layout(lib, "ref_dgr_image_layout", H * W, 2)
    res = {}
    g = out["geom"]
    res["depths"] = _slice(g, g_off[0], 4 * P, torch.float32, (P,))
    res["clamped"] = _slice(g, g_off[1], 3 * P, torch.uint8, (P, 3))
    res["means2D"] = _slice(g, g_off[3], 8 * P, torch.float32, (P, 2))
    res["cov3D"] = _slice(g, g_off[4], 24 * P, torch.float32, (P, 6))
    res["conic_opacity"] = _slice(g, g_off[5], 16 * P, torch.float32, (P, 4))
    res["rgb"] = _slice(g, g_off[6], 12 * P, torch.float32, (P, 3))
    res["tiles_touched"] = _slice(g, g_off[7], 4 * P, torch.int32, (P,))
    res["point_offsets"] = _slice(g, g_off[9], 4 * P, torch.int32, (P,))
    im = out["image"]
    res["n_contrib"] = _slice(im, i_off[0], 4 * H * W, torch.int32, (H, W))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    res["ranges"] = _slice(im, i_off[1], 8 * tiles, torch.int32, (tiles, 2))
    if N > 0:
        b_off, _ = _layout(lib, "ref_dgr_binning_layout", N, 5)
        b = out["binning"]
        res["point_list"] = _slice(b, b_off[0], 4 * N, torch.int32, (N,))
        res["point_list_keys"] = _slice(b, b_off[2], 8 * N, torch.int64, (N,))
    else:
        res["point_list"] = torch.zeros(0, dtype=torch.int32, device=g.device)
        res["point_list_keys"] = torch.zeros(0, dtype=torch.int64, device=g.device)
    return res


def dgr_backward(fwd, bg, means3D, colors, scales, rotations, scale_modifier, cov3D, view, proj, tanfovx,
                 tanfovy, dL_dcolor, dL_ddepth, dL_dalpha, sh, degree, campos, debug=False, sync=True):
    """Reference Rasterizer::backward with the ten zero-filled gradient tensors of the torch glue."""
    lib = _load("dgr")
    lib.ref_dgr_backward.restype = C.c_int
    dev = means3D.device
    P, M, H, W = fwd["P"], fwd["M"], fwd["H"], fwd["W"]
    z = lambda *s: torch.zeros(*s, device=dev)
    g = dict(means2D=z(P, 3), conic=z(P, 2, 2), opacity=z(P, 1), colors=z(P, 3), depths=z(P, 1), means3D=z(P, 3),
             cov3D=z(P, 6), sh=z(P, M, 3), scales=z(P, 3), rotations=z(P, 4))
    if sync:
        torch.cuda.synchronize(dev)
    rc = lib.ref_dgr_backward(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(fwd["num_rendered"]), _p(bg),
                              C.c_int(W), C.c_int(H), _p(means3D), _p(sh), _p(colors), _p(fwd["alpha"]), _p(scales),
                              C.c_float(scale_modifier), _p(rotations), _p(cov3D), _p(view), _p(proj), _p(campos),
                              C.c_float(tanfovx), C.c_float(tanfovy), _p(fwd["radii"]),
                              _aligned(fwd["geom"]), _aligned(fwd["binning"]), _aligned(fwd["image"]),
                              _p(dL_dcolor), _p(dL_ddepth), _p(dL_dalpha), _p(g["means2D"]), _p(g["conic"]),
                              _p(g["opacity"]), _p(g["colors"]), _p(g["depths"]), _p(g["means3D"]), _p(g["cov3D"]),
                              _p(g["sh"]), _p(g["scales"]), _p(g["rotations"]), C.c_int(int(debug)))
    if sync:
        torch.cuda.synchronize(dev)
    if rc != 0:
        raise RuntimeError("reference backward raised")
    return g


def _aligned(buf):
    """The reference re-derives sub-array pointers from the raw buffer pointer; pass it unchanged."""
    return None if buf is None else C.c_void_p(buf.data_ptr())


def dgr_mark_visible(means3D, view, proj):
    lib = _load("dgr")
    P = means3D.shape[0]
    present = torch.zeros(P, dtype=torch.bool, device=means3D.device)
    torch.cuda.synchronize()
    lib.ref_dgr_mark_visible(C.c_int(P), _p(means3D), _p(view), _p(proj), _p(present))
    torch.cuda.synchronize()
    return present


# ---- 2DGS surfels (diff-surfel-rasterization) --------------------------------------------------

def dsr_forward(bg, means3D, colors, opacity, scales, rotations, scale_modifier, transMat, view, proj,
                tanfovx, tanfovy, H, W, sh, degree, campos, prefiltered=False, debug=False, sync=True):
    """Reference surfel Rasterizer::forward (rasterizer_impl.cu:197-327)."""
    lib = _load("dsr")
    lib.ref_dsr_forward.restype = C.c_int
    dev = means3D.device
    P = means3D.shape[0]
    M = sh.shape[1] if sh is not None and sh.numel() else 0
    color = torch.zeros(3, H, W, device=dev); others = torch.zeros(7, H, W, device=dev)
    radii = torch.zeros(P, dtype=torch.int32, device=dev)
    ar = _Arena(dev)
    if sync:
        torch.cuda.synchronize(dev)
    n = lib.ref_dsr_forward(ar.fn, None, C.c_int(P), C.c_int(degree), C.c_int(M), _p(bg), C.c_int(W), C.c_int(H),
                            _p(means3D), _p(sh), _p(colors), _p(opacity), _p(scales), C.c_float(scale_modifier),
                            _p(rotations), _p(transMat), _p(view), _p(proj), _p(campos), C.c_float(tanfovx),
                            C.c_float(tanfovy), C.c_int(int(prefiltered)), _p(color), _p(others), _p(radii),
                            C.c_int(int(debug)))
    if sync:
        torch.cuda.synchronize(dev)
    if n < 0:
        raise RuntimeError("reference surfel forward raised")
    return dict(num_rendered=n, color=color, others=others, radii=radii,
                geom=ar.bufs.get(0), binning=ar.bufs.get(1), image=ar.bufs.get(2), P=P, M=M, H=H, W=W)


def dsr_intermediates(out):
    """Slice the reference's surfel scratch buffers (rasterizer_impl.h:21-72) into named tensors."""
    lib = _load("dsr")
    P, N, H, W = out["P"], out["num_rendered"], out["H"], out["W"]
    g_off, _ = _layout(lib, "ref_dsr_geom_layout", P, 10)
    i_off, _ = _layout(lib, "ref_dsr_image_layout", H * W, 3)
    res = {}
    g = out["geom"]
    res["depths"] = _slice(g, g_off[0], 4 * P, torch.float32, (P,))
    res["clamped"] = _slice(g, g_off[1], 3 * P, torch.uint8, (P, 3))
    res["means2D"] = _slice(g, g_off[3], 8 * P, torch.float32, (P, 2))
    res["transMat"] = _slice(g, g_off[4], 36 * P, torch.float32, (P, 9))
    res["normal_opacity"] = _slice(g, g_off[5], 16 * P, torch.float32, (P, 4))
    res["rgb"] = _slice(g, g_off[6], 12 * P, torch.float32, (P, 3))
    res["tiles_touched"] = _slice(g, g_off[7], 4 * P, torch.int32, (P,))
    im = out["image"]
    res["accum"] = _slice(im, i_off[0], 12 * H * W, torch.float32, (3, H, W))      # final_T, M1, M2
    res["n_contrib"] = _slice(im, i_off[1], 8 * H * W, torch.int32, (2, H, W))     # last, median
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    res["ranges"] = _slice(im, i_off[2], 8 * tiles, torch.int32, (tiles, 2))
    if N > 0:
        b_off, _ = _layout(lib, "ref_dsr_binning_layout", N, 5)
        b = out["binning"]
        res["point_list"] = _slice(b, b_off[0], 4 * N, torch.int32, (N,))
        res["point_list_keys"] = _slice(b, b_off[2], 8 * N, torch.int64, (N,))
    else:
        res["point_list"] = torch.zeros(0, dtype=torch.int32, device=g.device)
        res["point_list_keys"] = torch.zeros(0, dtype=torch.int64, device=g.device)
    return res


def dsr_backward(fwd, bg, means3D, colors, scales, rotations, scale_modifier, transMat, view, proj, tanfovx,
                 tanfovy, dL_dcolor, dL_dothers, sh, degree, campos, debug=False, sync=True):
    """Reference surfel Rasterizer::backward with the nine zero-filled tensors of the torch glue
    (rasterize_points.cu:192-200)."""
    lib = _load("dsr")
    lib.ref_dsr_backward.restype = C.c_int
    dev = means3D.device
    P, M, H, W = fwd["P"], fwd["M"], fwd["H"], fwd["W"]
    z = lambda *s: torch.zeros(*s, device=dev)
    g = dict(means2D=z(P, 3), normal=z(P, 3), opacity=z(P, 1), colors=z(P, 3), means3D=z(P, 3), transMat=z(P, 9),
             sh=z(P, M, 3), scales=z(P, 2), rotations=z(P, 4))
    if sync:
        torch.cuda.synchronize(dev)
    rc = lib.ref_dsr_backward(C.c_int(P), C.c_int(degree), C.c_int(M), C.c_int(fwd["num_rendered"]), _p(bg),
                              C.c_int(W), C.c_int(H), _p(means3D), _p(sh), _p(colors), _p(scales),
                              C.c_float(scale_modifier), _p(rotations), _p(transMat), _p(view), _p(proj), _p(campos),
                              C.c_float(tanfovx), C.c_float(tanfovy), _p(fwd["radii"]),
                              _aligned(fwd["geom"]), _aligned(fwd["binning"]), _aligned(fwd["image"]),
                              _p(dL_dcolor), _p(dL_dothers), _p(g["means2D"]), _p(g["normal"]), _p(g["opacity"]),
                              _p(g["colors"]), _p(g["means3D"]), _p(g["transMat"]), _p(g["sh"]), _p(g["scales"]),
                              _p(g["rotations"]), C.c_int(int(debug)))
    if sync:
        torch.cuda.synchronize(dev)
    if rc != 0:
        raise RuntimeError("reference surfel backward raised")
    return g


# ---- simple-knn -----------------------------------------------------------------------------------

def knn_dist2(points):
    """Reference SimpleKNN::knn (simple_knn.cu:186-221) behind distCUDA2's allocation (spatial.cu:15-26)."""
    lib = _load("knn")
    pts = points.contiguous().float()
    P = pts.shape[0]
    out = torch.zeros(P, dtype=torch.float32, device=pts.device)
    torch.cuda.synchronize()
    lib.ref_knn_dist2(C.c_int(P), _p(pts), _p(out))
    torch.cuda.synchronize()
    return out
