"""ctypes binding of libgsr_b200.so -- the C ABI declared in include/gsr_b200.h.

The library is plain CUDA C++ (no torch headers); this module only declares signatures and
hands over raw device pointers of torch tensors. There is NO fallback: if the shared object
is missing or a symbol is absent, importing the product packages raises.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libgsr_b200.so"

c_void_p, c_int, c_size_t, c_float = C.c_void_p, C.c_int, C.c_size_t, C.c_float

READY_FN = C.CFUNCTYPE(None, C.c_void_p)          # gsr_backward_early_rgb callback
CHUNK_FN = C.CFUNCTYPE(None, C.c_void_p, C.c_int, C.c_int, C.c_int)     # gsr_backward_peer_head: chunk_done(user, k, row0, rows)
ALLOC_FN = C.CFUNCTYPE(c_void_p, c_void_p, c_int, c_size_t)

GSR_BUF_GEOM, GSR_BUF_BINNING, GSR_BUF_IMAGE = 0, 1, 2


class GsrSettings(C.Structure):
    """struct GsrSettings (include/gsr_b200.h)."""
    _fields_ = [
        ("image_height", C.c_int32), ("image_width", C.c_int32),
        ("tanfovx", c_float), ("tanfovy", c_float), ("scale_modifier", c_float),
        ("sh_degree", C.c_int32), ("prefiltered", C.c_int32), ("debug", C.c_int32),
        ("bg", c_void_p), ("viewmatrix", c_void_p), ("projmatrix", c_void_p), ("campos", c_void_p),
    ]


class GsrModelParams(C.Structure):
    """struct GsrModelParams (include/gsr_b200.h): the tensors GaussianModel stores."""
    _fields_ = [(n, c_void_p) for n in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")]


class GsrModelGrads(C.Structure):
    """struct GsrModelGrads (include/gsr_b200.h)."""
    _fields_ = [(n, c_void_p) for n in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "rotation")]


class GsrAdamGroup(C.Structure):
    """struct GsrAdamGroup (include/gsr_b200.h)."""
    _fields_ = [("param", c_void_p), ("grad", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p),
                ("n", C.c_int64), ("lr", C.c_double)]


class GsrDensifyCfg(C.Structure):
    """struct GsrDensifyCfg (include/gsr_b200.h)."""
    _fields_ = [("grad_threshold", c_float), ("dense_extent", c_float), ("min_opacity", c_float),
                ("big_extent", c_float), ("max_screen_size", c_float), ("prune_big", c_int), ("n_split", c_int)]


class GsrDensifyGroup(C.Structure):
    """struct GsrDensifyGroup (include/gsr_b200.h)."""
    _fields_ = [("param", c_void_p), ("exp_avg", c_void_p), ("exp_avg_sq", c_void_p), ("out_param", c_void_p),
                ("out_exp_avg", c_void_p), ("out_exp_avg_sq", c_void_p), ("width", c_int), ("kind", c_int)]


_P = c_void_p
# name -> (restype, argtypes); every symbol include/gsr_b200.h declares
SIGNATURES = {
    "gsr_last_error": (C.c_char_p, []),
    "gsr_version": (c_int, []),
    "gsr_profile_enable": (None, [c_int]),
    "gsr_profile_read": (c_int, [C.POINTER(c_float)]),
    "gsr_rasterize_forward": (c_int, [C.POINTER(GsrSettings), c_int, c_int] + [_P] * 7 + [_P] * 4
                              + [ALLOC_FN, _P, C.POINTER(c_int), _P]),
    "gsr_rasterize_forward_nosync": (c_int, [C.POINTER(GsrSettings), c_int, c_int] + [_P] * 7 + [_P] * 4
                                     + [ALLOC_FN, _P, c_int, _P, _P]),
    "gsr_surfel_forward_nosync": (c_int, [C.POINTER(GsrSettings), c_int, c_int] + [_P] * 7 + [_P] * 3
                                  + [ALLOC_FN, _P, c_int, _P, _P]),
    "gsr_backward_workspace_bytes": (c_size_t, [c_int]),
    "gsr_rasterize_backward": (c_int, [C.POINTER(GsrSettings), c_int, c_int, c_int] + [_P] * 6 + [_P]
                               + [_P] * 3 + [_P] + [_P] * 3 + [_P] * 8 + [_P, c_size_t, _P]),
    "gsr_rasterize_forward_raw": (c_int, [C.POINTER(GsrSettings), c_int, c_int, C.POINTER(GsrModelParams)] + [_P] * 4
                                  + [ALLOC_FN, _P, C.POINTER(c_int), _P]),
    "gsr_rasterize_backward_raw": (c_int, [C.POINTER(GsrSettings), c_int, c_int, c_int, C.POINTER(GsrModelParams)]
                                   + [_P] * 5 + [_P] * 3 + [_P, C.POINTER(GsrModelGrads), _P, c_size_t, _P]),
    "gsr_surfel_forward_raw": (c_int, [C.POINTER(GsrSettings), c_int, c_int, C.POINTER(GsrModelParams)] + [_P] * 3
                               + [ALLOC_FN, _P, C.POINTER(c_int), _P]),
    "gsr_surfel_backward_raw": (c_int, [C.POINTER(GsrSettings), c_int, c_int, c_int, C.POINTER(GsrModelParams)]
                                + [_P] * 4 + [_P] * 2 + [_P, C.POINTER(GsrModelGrads), _P, c_size_t, _P]),
    "gsr_rasterize_forward_views": (c_int, [C.POINTER(GsrSettings), c_int, c_int, c_int] + [_P] * 7 + [_P] * 4
                                    + [ALLOC_FN, _P, C.POINTER(c_int), _P]),
    "gsr_adam_step": (c_int, [c_int, C.POINTER(GsrAdamGroup), C.c_double, C.c_double, C.c_double, c_int, _P]),
    "gsr_densification_stats": (c_int, [c_int, _P, _P, c_int, _P, _P, _P, _P]),
    "gsr_reset_opacity": (c_int, [c_int, _P, _P, _P, c_float, _P]),
    "gsr_densify_workspace_bytes": (c_size_t, [c_int, c_int]),
    "gsr_densify_plan": (c_int, [c_int, c_int, _P, _P, _P, _P, C.POINTER(GsrDensifyCfg), _P, C.POINTER(C.c_int32), _P]),
    "gsr_densify_apply": (c_int, [c_int, c_int, C.POINTER(GsrDensifyCfg), _P, C.POINTER(C.c_int32), c_int,
                                  C.POINTER(GsrDensifyGroup), _P, _P, _P, _P]),
    "gsr_ply_row_floats": (c_int, [c_int, c_int]),
    "gsr_ply_pack": (c_int, [c_int, c_int, c_int, C.POINTER(GsrModelParams), _P, _P]),
    "gsr_ply_unpack": (c_int, [c_int, c_int, _P, C.POINTER(C.c_int32), c_int, c_int, C.POINTER(GsrModelGrads), _P]),
    "gsr_l1_ssim_workspace_bytes": (c_size_t, [c_int, c_int, c_int]),
    "gsr_l1_ssim_forward": (c_int, [c_int, c_int, c_int, _P, _P, c_float, _P, _P, c_size_t, _P]),
    "gsr_l1_ssim_backward": (c_int, [c_int, c_int, c_int, _P, _P, _P, c_float, c_float, _P, _P, _P]),
    "gsr_mark_visible": (c_int, [c_int, _P, _P, _P, _P, _P]),
    "gsr_geom_layout": (None, [c_int, c_int, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "gsr_image_layout": (None, [c_int, c_int, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "gsr_binning_layout": (None, [c_int, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "gsr_surfel_forward": (c_int, [C.POINTER(GsrSettings), c_int, c_int] + [_P] * 7 + [_P] * 3
                           + [ALLOC_FN, _P, C.POINTER(c_int), _P]),
    "gsr_surfel_backward_workspace_bytes": (c_size_t, [c_int]),
    "gsr_surfel_backward": (c_int, [C.POINTER(GsrSettings), c_int, c_int, c_int] + [_P] * 6 + [_P]
                            + [_P] * 3 + [_P] * 2 + [_P] * 8 + [_P, c_size_t, _P]),
    "gsr_surfel_geom_layout": (None, [c_int, c_int, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "gsr_surfel_image_layout": (None, [c_int, c_int, C.POINTER(c_size_t), C.POINTER(c_size_t)]),
    "gsr_knn_workspace_bytes": (c_size_t, [c_int]),
    "gsr_knn_dist2": (c_int, [c_int, _P, _P, _P, c_size_t, _P]),
    "gsr_knn_static_max_points": (c_int, []),
    "gsr_knn_dist2_static": (c_int, [c_int, _P, _P, _P]),
    "gsr_clamped_color_grad": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P]),
    "gsr_backward_early_rgb": (c_int, [_P, _P, _P]),
    "gsr_backward_early_rgb_peers": (c_int, [c_int, C.POINTER(c_void_p), C.POINTER(c_void_p), _P, _P, _P, _P]),
    "gsr_backward_early_rgb_multicast": (c_int, [_P, _P, _P, _P, _P, _P]),
    "gsr_backward_peer_head": (c_int, [_P, c_int, _P, _P]),
    "gsr_peer_consume": (c_int, [_P, _P, c_int, c_int, c_int, c_int, _P]),
    "gsr_peer_reduce_shard": (c_int, [_P, c_int, c_int, c_int, c_int, _P]),
    "gsr_peer_finish": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P, _P, c_int, _P]),
    "gsr_sh_grad_from_views": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P]),
}

_lib = None


class GsrError(RuntimeError):
    pass


def lib() -> C.CDLL:
    """Load (once) and return the shared library; raises loudly if it is not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise GsrError(
            f"{LIB_PATH} is missing: build it with `python -m scorp_b200.build` "
            "(or __graft_entry__.build()). There is no CPU or PyTorch fallback.")
    handle = C.CDLL(str(LIB_PATH))
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(handle, name)          # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = handle
    return _lib


def check(rc: int, what: str) -> None:
    if rc != 0:
        msg = lib().gsr_last_error()
        raise GsrError(f"{what} failed (code {rc}): {msg.decode() if msg else '?'}")


def ptr(t):
    """Device pointer of a tensor, or None (NULL) for the reference's 0-element sentinel."""
    if t is None or t.numel() == 0:
        return None
    return t.data_ptr()
