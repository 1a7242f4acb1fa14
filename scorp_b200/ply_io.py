"""The model's on-disk format: GaussianModel.save_ply / load_ply without plyfile and without per-Gaussian host loops
(last item).

Reference: gs3dgs/scene/gaussian_model.py:220-251 (construct_list_of_attributes, save_ply) and :287-332 (load_ply);
gs2dgs/scene/gaussian_model.py has the same pair with two scales. The file is a binary little-endian PLY with one
`vertex` element of float properties  x y z nx ny nz f_dc_* f_rest_* opacity scale_* rot_*  -- what
`PlyData([PlyElement.describe(elements, 'vertex')]).write(path)` produces (plyfile is a third-party dependency of the
reference that is not in this image; the header below follows the PLY specification and plyfile's naming of 'f4' as
`float`, and the reader accepts both `float` and `float32`).

The vertex rows are packed and unpacked on the GPU (csrc/ply.cu), so the model crosses PCIe in one copy through pinned
memory; the reference goes tensor by tensor and builds the structured array with `list(map(tuple, attributes))`.
CUDA tensors only. Files with non-float properties or ascii / big-endian encoding are converted column by column on the
host first (numpy), then take the same path.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Tuple

import numpy as np
import torch

from . import _lib

__all__ = ["construct_list_of_attributes", "save_ply", "load_ply", "load_ply_into"]

_ATTRS = (("xyz", "_xyz"), ("features_dc", "_features_dc"), ("features_rest", "_features_rest"), ("opacity", "_opacity"),
          ("scaling", "_scaling"), ("rotation", "_rotation"))
_NP_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2", "ushort": "u2",
             "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4", "float": "f4", "float32": "f4",
             "double": "f8", "float64": "f8"}


def construct_list_of_attributes(sh_rest: int, scale_dim: int = 3) -> List[str]:
    """gaussian_model.py:220-232 for a model with `sh_rest` coefficients per channel besides DC."""
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += [f"f_dc_{i}" for i in range(3)]
    names += [f"f_rest_{i}" for i in range(3 * sh_rest)]
    names.append("opacity")
    names += [f"scale_{i}" for i in range(scale_dim)]
    names += [f"rot_{i}" for i in range(4)]
    return names


def _tensors(model) -> Dict[str, torch.Tensor]:
    if isinstance(model, dict):
        return {k: model[k] for k, _ in _ATTRS}
    return {k: getattr(model, a) for k, a in _ATTRS}


def _stream(dev):
    return C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)


def save_ply(model, path: str) -> None:
    """GaussianModel.save_ply(path) (gaussian_model.py:234-251). `model` is a GaussianModel-shaped object (`_xyz`,
    `_features_dc` [P,1,3], `_features_rest` [P,R,3], `_opacity` [P,1], `_scaling` [P,S], `_rotation` [P,4]) or a dict
    with the keys xyz, features_dc, features_rest, opacity, scaling, rotation."""
    lib = _lib.lib()
    t = {k: v.detach() for k, v in _tensors(model).items()}
    xyz = t["xyz"]
    if not xyz.is_cuda:
        raise _lib.GsrError("save_ply packs the vertex rows on the GPU: CUDA tensors only, there is no CPU path")
    P, dev = xyz.shape[0], xyz.device
    for k, v in t.items():
        if v.dtype != torch.float32 or v.shape[0] != P or v.device != dev:
            raise RuntimeError(f"{k}: expected an fp32 tensor with {P} rows on {dev}")
        t[k] = v.contiguous()
    if t["features_dc"].shape[1:] != (1, 3) or t["features_rest"].dim() != 3 or t["features_rest"].shape[2] != 3:
        raise RuntimeError("expected features_dc [P,1,3] and features_rest [P,R,3]")
    R, S = t["features_rest"].shape[1], t["scaling"].shape[1]
    names = construct_list_of_attributes(R, S)
    assert len(names) == int(lib.gsr_ply_row_floats(R, S))
    header = "ply\nformat binary_little_endian 1.0\n" + f"element vertex {P}\n" \
             + "".join(f"property float {n}\n" for n in names) + "end_header\n"
    host = torch.empty((P, len(names)), dtype=torch.float32, pin_memory=True)
    if P:
        rows = torch.empty((P, len(names)), device=dev)
        mp = _lib.GsrModelParams(*[_lib.ptr(t[k]) for k, _ in _ATTRS])
        with torch.cuda.device(dev):
            _lib.check(lib.gsr_ply_pack(P, R, S, C.byref(mp), _lib.ptr(rows), _stream(dev)), "gsr_ply_pack")
            host.copy_(rows, non_blocking=True)
            torch.cuda.current_stream(dev).synchronize()
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)                  # mkdir_p(os.path.dirname(path)), :235
    with open(path, "wb") as f:
        f.write(header.encode("ascii"))
        if P:
            f.write(memoryview(host.numpy()).cast("B"))


def _read_header(f) -> Tuple[str, int, List[Tuple[str, str]], List[str]]:
    """-> (format, vertex count, [(name, numpy type)] of the first element, names of later elements)"""
    if f.readline().strip() != b"ply":
        raise ValueError("not a PLY file")
    fmt, count, props, later, current = None, 0, [], [], None
    while True:
        line = f.readline()
        if not line:
            raise ValueError("PLY header without end_header")
        w = line.decode("ascii", "replace").split()
        if not w or w[0] in ("comment", "obj_info"):
            continue
        if w[0] == "format":
            fmt = w[1]
        elif w[0] == "element":
            if current is None:
                current, count = w[1], int(w[2])
            else:
                current = w[1]; later.append(w[1])
        elif w[0] == "property":
            if w[1] == "list":
                if not later:
                    raise ValueError("list properties in the first PLY element are not a Gaussian model")
                continue
            if not later:
                if w[1] not in _NP_TYPES:
                    raise ValueError(f"unknown PLY property type {w[1]}")
                props.append((w[2], _NP_TYPES[w[1]]))
        elif w[0] == "end_header":
            break
    if fmt is None or current is None:
        raise ValueError("PLY header lacks a format or an element")
    return fmt, count, props, later


def _rows_from_file(path: str) -> Tuple[torch.Tensor, List[str]]:
    """-> (pinned float32 [P, n_props] rows of the first element, property names)"""
    with open(path, "rb") as f:
        fmt, P, props, _ = _read_header(f)
        names = [n for n, _ in props]
        host = torch.empty((P, len(props)), dtype=torch.float32, pin_memory=torch.cuda.is_available())
        dst = host.numpy()
        if P == 0 or not props:
            pass
        elif fmt == "ascii":
            for i in range(P):
                dst[i] = np.array(f.readline().split()[:len(props)], dtype=np.float64)
        else:
            order = "<" if fmt == "binary_little_endian" else ">"
            if all(ty == "f4" for _, ty in props) and order == "<":
                n = f.readinto(memoryview(dst).cast("B"))      # straight into pinned memory
                if n != dst.nbytes:
                    raise ValueError("PLY vertex data is truncated")
            else:                                              # mixed types: one numpy pass over the columns
                rec = np.fromfile(f, dtype=np.dtype([(n, order + ty) for n, ty in props]), count=P)
                if rec.shape[0] != P:
                    raise ValueError("PLY vertex data is truncated")
                for c, n in enumerate(names):
                    dst[:, c] = rec[n]
    return host, names


def load_ply(path: str, max_sh_degree: int = None, device="cuda") -> Dict[str, torch.Tensor]:
    """The tensors GaussianModel.load_ply builds (gaussian_model.py:287-327) as a dict of CUDA fp32 tensors:
    xyz [P,3], features_dc [P,1,3], features_rest [P,R,3], opacity [P,1], scaling [P,S], rotation [P,4].
    With max_sh_degree the reference's assert on the number of f_rest_* properties (:304) is applied."""
    lib = _lib.lib()
    dev = torch.device(device)
    if dev.type != "cuda":
        raise _lib.GsrError("load_ply unpacks the vertex rows on the GPU: a CUDA device is required, there is no CPU path")
    if dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    host, names = _rows_from_file(path)
    col = {n: i for i, n in enumerate(names)}
    by_index = lambda prefix: sorted((n for n in names if n.startswith(prefix)), key=lambda x: int(x.split("_")[-1]))
    rest, scales, rots = by_index("f_rest_"), by_index("scale_"), by_index("rot")
    if max_sh_degree is not None:
        assert len(rest) == 3 * (max_sh_degree + 1) ** 2 - 3
    if len(rest) % 3 or len(rots) != 4 or not scales:
        raise ValueError("PLY file is not a Gaussian model: f_rest_* must be a multiple of 3, rot_* four, scale_* present")
    R, S, P = len(rest) // 3, len(scales), host.shape[0]
    try:
        columns = [col["x"], col["y"], col["z"], col["f_dc_0"], col["f_dc_1"], col["f_dc_2"]]
        columns += [col[rest[c * R + k]] for k in range(R) for c in range(3)]       # features_rest[i][k][c] = f_rest_{c*R+k}
        columns += [col["opacity"]] + [col[n] for n in scales] + [col[n] for n in rots]
    except KeyError as e:
        raise ValueError(f"PLY file lacks the property {e}") from None
    out = dict(xyz=torch.empty((P, 3), device=dev), features_dc=torch.empty((P, 1, 3), device=dev),
               features_rest=torch.empty((P, R, 3), device=dev), opacity=torch.empty((P, 1), device=dev),
               scaling=torch.empty((P, S), device=dev), rotation=torch.empty((P, 4), device=dev))
    if P:
        with torch.cuda.device(dev):
            rows = host.to(dev, non_blocking=True)
            arr = (C.c_int32 * len(columns))(*columns)
            mg = _lib.GsrModelGrads(*[_lib.ptr(out[k]) for k, _ in _ATTRS])
            _lib.check(lib.gsr_ply_unpack(P, host.shape[1], _lib.ptr(rows), arr, R, S, C.byref(mg), _stream(dev)),
                       "gsr_ply_unpack")
            torch.cuda.current_stream(dev).synchronize()          # the pinned staging buffer is released on return
    return out


def load_ply_into(gaussians, path: str, spatial_lr_scale=None) -> None:
    """GaussianModel.load_ply(path, spatial_lr_scale) on a GaussianModel-shaped object (gaussian_model.py:287-332): sets the
    six nn.Parameters, active_sh_degree = max_sh_degree and a zeroed max_radii2D."""
    gaussians.spatial_lr_scale = spatial_lr_scale
    t = load_ply(path, max_sh_degree=gaussians.max_sh_degree)
    for k, attr in _ATTRS:
        setattr(gaussians, attr, torch.nn.Parameter(t[k].requires_grad_(True)))
    gaussians.active_sh_degree = gaussians.max_sh_degree
    gaussians.max_radii2D = torch.zeros((t["xyz"].shape[0]), device=t["xyz"].device)
