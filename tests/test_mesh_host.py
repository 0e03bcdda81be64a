"""CPU: the host-side mesh container that stands in for trimesh (reference requirements.txt:34) on the extraction
path -- face-adjacency components (trimesh.Trimesh.split, reference mesh.py:151-158), concatenation and the PLY
writer (scripts/extract_mesh.py:92-101 exports through trimesh)."""
import struct

import numpy as np

from neuralangelo_b200 import mesh as na_mesh


def _tetra(offset):
    v = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [0, 0, 1]], dtype=np.float64) + offset
    f = np.array([[0, 2, 1], [0, 1, 3], [0, 3, 2], [1, 2, 3]], dtype=np.int64)
    return v, f


def test_split_uses_shared_edges_not_shared_vertices():
    """Two tetrahedra welded in ONE vertex are two components for trimesh (face adjacency = shared edge);
    a vertex-based union-find would call them one."""
    v0, f0 = _tetra(0.0)
    v1, f1 = _tetra(0.0)
    v1 = -2.0 * v1                       # bigger, mirrored through the origin: touches v0 only in vertex 0
    verts = np.concatenate([v0, v1[1:]])                  # share vertex 0
    remap = np.array([0, 4, 5, 6])
    faces = np.concatenate([f0, remap[f1]])
    m = na_mesh.Mesh(verts, faces)
    comps = m.split()
    assert len(comps) == 2
    assert sorted(len(c.faces) for c in comps) == [4, 4]
    lcc = na_mesh.filter_largest_cc(m)
    assert len(lcc.faces) == 4 and np.isclose(lcc.area, 4.0 * na_mesh.Mesh(v0, f0).area)
    # faces that share an edge stay together
    labels = na_mesh._face_components(f0, 4)
    assert len(np.unique(labels)) == 1
    assert len(na_mesh._face_components(np.zeros((0, 3), dtype=np.int64), 0)) == 0


def test_concatenate_offsets_faces():
    a = na_mesh.Mesh(*_tetra(0.0))
    b = na_mesh.Mesh(*_tetra(5.0))
    c = na_mesh.concatenate([a, b])
    assert c.vertices.shape == (8, 3) and c.faces.shape == (8, 3)
    assert c.faces[4:].min() == 4 and c.faces[4:].max() == 7
    assert len(c.split()) == 2


def test_ply_export_round_trip(tmp_path):
    v, f = _tetra(0.25)
    colors = np.arange(12, dtype=np.uint8).reshape(4, 3) * 20
    m = na_mesh.Mesh(v, f, colors)
    path = tmp_path / "m.ply"
    m.export(str(path))
    raw = path.read_bytes()
    head, body = raw.split(b"end_header\n", 1)
    lines = head.decode("ascii").splitlines()
    assert lines[0] == "ply" and lines[1] == "format binary_little_endian 1.0"
    assert "element vertex 4" in lines and "element face 4" in lines
    assert "property list uchar int vertex_indices" in lines
    assert len(body) == 4 * (12 + 4) + 4 * (1 + 12)
    for i in range(4):
        x, y, z, r, g, b, a = struct.unpack_from("<fffBBBB", body, i * 16)
        np.testing.assert_allclose([x, y, z], v[i].astype(np.float32))
        assert (r, g, b, a) == (*colors[i], 255)
    for i in range(4):
        n, i0, i1, i2 = struct.unpack_from("<Biii", body, 64 + i * 13)
        assert n == 3 and [i0, i1, i2] == list(f[i])
