"""TEST INFRASTRUCTURE (checker): numpy restatement of GaussianModel.save_ply / load_ply. Never imported by scorp_b200/.

Follows /root/reference/gs3dgs/scene/gaussian_model.py:220-251 (construct_list_of_attributes, save_ply) and :287-327
(load_ply) statement by statement; gs2dgs/scene/gaussian_model.py has the same pair with two scale columns.
The byte format itself is written by plyfile (`PlyData([PlyElement.describe(elements, 'vertex')]).write(path)`), a
third-party dependency of the reference that is absent from /root/reference and from this image (no pinned version in
the reference: requirements list it unversioned). Restated from the PLY specification: header lines `ply`, `format
binary_little_endian 1.0`, `element vertex N`, one `property float <name>` per 'f4' field, `end_header`, then the
structured array's bytes. Parity status: the ROW CONTENT follows the reference's own numpy code; the HEADER BYTES are
unpinned (plyfile cannot be run here)."""
from __future__ import annotations

import numpy as np

f32 = np.float32


def construct_list_of_attributes(features_dc, features_rest, scaling, rotation):
    """:220-232"""
    l = ['x', 'y', 'z', 'nx', 'ny', 'nz']
    for i in range(features_dc.shape[1] * features_dc.shape[2]):
        l.append('f_dc_{}'.format(i))
    for i in range(features_rest.shape[1] * features_rest.shape[2]):
        l.append('f_rest_{}'.format(i))
    l.append('opacity')
    for i in range(scaling.shape[1]):
        l.append('scale_{}'.format(i))
    for i in range(rotation.shape[1]):
        l.append('rot_{}'.format(i))
    return l


def vertex_rows(m):
    """:236-247 -> (names, float32 [P, n]); m has xyz, features_dc [P,1,3], features_rest [P,R,3], opacity, scaling, rotation"""
    xyz = np.asarray(m["xyz"], f32)
    normals = np.zeros_like(xyz)
    flat = lambda t: t.transpose(0, 2, 1).reshape(t.shape[0], t.shape[1] * t.shape[2])     # transpose(1, 2).flatten(start_dim=1)
    f_dc = flat(np.asarray(m["features_dc"], f32))
    f_rest = flat(np.asarray(m["features_rest"], f32))
    names = construct_list_of_attributes(m["features_dc"], m["features_rest"], m["scaling"], m["rotation"])
    rows = np.concatenate((xyz, normals, f_dc, f_rest, np.asarray(m["opacity"], f32), np.asarray(m["scaling"], f32),
                           np.asarray(m["rotation"], f32)), axis=1)
    return names, rows


def write_ply(path, names, rows, types=None, byte_order="<", comments=()):
    """the structured array + header; `types` maps a property name to a numpy type other than f4 (reader tests)"""
    types = types or {}
    ply_name = {"f4": "float", "f8": "double", "u1": "uchar", "i4": "int", "i2": "short"}
    dt = np.dtype([(n, byte_order + types.get(n, "f4")) for n in names])
    el = np.empty(rows.shape[0], dtype=dt)
    for c, n in enumerate(names):
        el[n] = rows[:, c]
    fmt = "binary_little_endian" if byte_order == "<" else "binary_big_endian"
    with open(path, "wb") as f:
        f.write(("ply\nformat %s 1.0\n" % fmt).encode())
        for c in comments:
            f.write(("comment %s\n" % c).encode())
        f.write(("element vertex %d\n" % rows.shape[0]).encode())
        for n in names:
            f.write(("property %s %s\n" % (ply_name[types.get(n, "f4")], n)).encode())
        f.write(b"end_header\n")
        el.tofile(f)


def model_from_columns(names, rows, max_sh_degree):
    """load_ply :292-326 on the columns of the vertex element -> the six tensors (float32, the reference's shapes)"""
    col = {n: np.asarray(rows[:, i]) for i, n in enumerate(names)}
    xyz = np.stack((col["x"], col["y"], col["z"]), axis=1)
    opacities = col["opacity"][..., np.newaxis]
    features_dc = np.zeros((xyz.shape[0], 3, 1))
    features_dc[:, 0, 0] = col["f_dc_0"]; features_dc[:, 1, 0] = col["f_dc_1"]; features_dc[:, 2, 0] = col["f_dc_2"]
    extra = sorted([n for n in names if n.startswith("f_rest_")], key=lambda x: int(x.split('_')[-1]))
    assert len(extra) == 3 * (max_sh_degree + 1) ** 2 - 3
    features_extra = np.zeros((xyz.shape[0], len(extra)))
    for idx, n in enumerate(extra):
        features_extra[:, idx] = col[n]
    features_extra = features_extra.reshape((features_extra.shape[0], 3, (max_sh_degree + 1) ** 2 - 1))
    scale_names = sorted([n for n in names if n.startswith("scale_")], key=lambda x: int(x.split('_')[-1]))
    scales = np.stack([col[n] for n in scale_names], axis=1)
    rot_names = sorted([n for n in names if n.startswith("rot")], key=lambda x: int(x.split('_')[-1]))
    rots = np.stack([col[n] for n in rot_names], axis=1)
    return dict(xyz=xyz.astype(f32), features_dc=features_dc.astype(f32).transpose(0, 2, 1).copy(),
                features_rest=features_extra.astype(f32).transpose(0, 2, 1).copy(), opacity=opacities.astype(f32),
                scaling=scales.astype(f32), rotation=rots.astype(f32))
