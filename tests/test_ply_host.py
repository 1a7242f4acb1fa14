"""PLY model files on the host side (no GPU): the numpy oracle's save/load restatement round-trips, and the product's
header parser and row reader (scorp_b200/ply_io.py: host logic, no compute) read what the oracle writes -- float,
mixed-type, big-endian and ascii files, comments, a second element, any property order."""
import numpy as np
import pytest
import torch

from oracle import ply_oracle as PO
from scorp_b200 import ply_io


def _model(P, R, S, seed=0):
    g = np.random.default_rng(seed)
    r = lambda *s: g.standard_normal(s).astype(np.float32)
    return dict(xyz=r(P, 3), features_dc=r(P, 1, 3), features_rest=r(P, R, 3), opacity=r(P, 1), scaling=r(P, S), rotation=r(P, 4))


@pytest.mark.parametrize("R,S,deg", [(15, 3, 3), (0, 3, 0), (3, 2, 1)])
def test_oracle_round_trip(R, S, deg):
    m = _model(37, R, S)
    names, rows = PO.vertex_rows(m)
    assert names == ply_io.construct_list_of_attributes(R, S) and rows.shape == (37, len(names))
    assert np.all(rows[:, 3:6] == 0)
    if R:   # channel-major SH: f_rest_{c*R+k} = features_rest[i][k][c]
        assert rows[5, names.index("f_rest_%d" % (1 * R + 2))] == m["features_rest"][5, 2, 1]
    back = PO.model_from_columns(names, rows, deg)
    for k in m:
        assert back[k].shape == m[k].shape and np.array_equal(back[k], m[k]), k


def test_row_reader_reads_every_encoding(tmp_path):
    m = _model(23, 3, 3, seed=4)
    names, rows = PO.vertex_rows(m)
    perm = list(np.random.default_rng(1).permutation(len(names)))
    pn, pr = [names[i] for i in perm], rows[:, perm]
    files = {}
    PO.write_ply(tmp_path / "le.ply", names, rows); files["le"] = (names, rows)
    PO.write_ply(tmp_path / "perm.ply", pn, pr, comments=("written by a test", "")); files["perm"] = (pn, pr)
    PO.write_ply(tmp_path / "mixed.ply", names, rows, types={"opacity": "f8", "nx": "u1", "ny": "i4"}); files["mixed"] = (names, rows)
    PO.write_ply(tmp_path / "be.ply", names, rows, byte_order=">"); files["be"] = (names, rows)
    with open(tmp_path / "ascii.ply", "w") as f:
        f.write("ply\nformat ascii 1.0\nelement vertex %d\n" % rows.shape[0] + "".join("property float32 %s\n" % n for n in names)
                + "element face 0\nproperty list uchar int vertex_indices\nend_header\n")
        for r in rows:
            f.write(" ".join(repr(float(x)) for x in r) + "\n")
    files["ascii"] = (names, rows)
    for key, (n_ref, r_ref) in files.items():
        host, got_names = ply_io._rows_from_file(str(tmp_path / f"{key}.ply"))
        assert got_names == n_ref, key
        assert host.dtype == torch.float32 and np.array_equal(host.numpy(), r_ref), key


def test_bad_files(tmp_path):
    p = tmp_path / "x.ply"
    p.write_bytes(b"plx\n")
    with pytest.raises(ValueError, match="not a PLY"):
        ply_io._rows_from_file(str(p))
    p.write_bytes(b"ply\nformat binary_little_endian 1.0\nelement vertex 3\nproperty float x\nend_header\n\0\0\0\0")
    with pytest.raises(ValueError, match="truncated"):
        ply_io._rows_from_file(str(p))
    p.write_bytes(b"ply\nformat binary_little_endian 1.0\nelement vertex 3\nproperty float x\n")
    with pytest.raises(ValueError, match="end_header"):
        ply_io._rows_from_file(str(p))


def test_no_cpu_path():
    from scorp_b200 import _lib
    m = {k: torch.from_numpy(v) for k, v in _model(4, 0, 3).items()}
    with pytest.raises(_lib.GsrError):
        ply_io.save_ply(m, "/tmp/never_written.ply")
    with pytest.raises(_lib.GsrError):
        ply_io.load_ply("/tmp/never_read.ply", device="cpu")
