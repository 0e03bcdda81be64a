"""CPU: the marching-cubes oracle (oracle/mcubes_ref.c) has no reference-produced golden vector
(PyMCubes is not installable here: parity unpinned), so it is anchored on analytic properties."""
import numpy as np
import pytest

from oracle import mcubes_ref


def sphere_volume(n, radius=0.5131, center=(0.013, -0.021, 0.007)):
    ax = np.linspace(-1, 1, n, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax, ax, indexing="ij")
    c = np.asarray(center, dtype=np.float32)
    return (np.sqrt((x - c[0]) ** 2 + (y - c[1]) ** 2 + (z - c[2]) ** 2) - radius).astype(np.float32), ax


def edge_counts(faces):
    e = np.concatenate([faces[:, [0, 1]], faces[:, [1, 2]], faces[:, [2, 0]]], axis=0)
    und = np.sort(e, axis=1)
    _, cnt = np.unique(und, axis=0, return_counts=True)
    return e, cnt


def test_single_cell_cases_match_table():
    """One cell, one corner inside (case 1 -> triangle on edges 0, 8, 3)."""
    vol = np.ones((2, 2, 2), dtype=np.float32)
    vol[0, 0, 0] = -1.0
    v, f, cases = mcubes_ref.marching_cubes(vol, 0.0, return_cases=True)
    assert cases[0, 0, 0] == 1 and f.shape == (1, 3) and v.shape == (3, 3)
    got = {tuple(np.round(p, 6)) for p in v}
    assert got == {(0.5, 0, 0), (0, 0.5, 0), (0, 0, 0.5)}
    # triangle 0 8 3 = (x-edge, z-edge, y-edge) vertices in that order
    np.testing.assert_allclose(v[f[0]], [[0.5, 0, 0], [0, 0, 0.5], [0, 0.5, 0]])


def test_sphere_is_closed_manifold_with_euler_2():
    vol, ax = sphere_volume(48)
    v, f = mcubes_ref.marching_cubes(vol, 0.0)
    assert len(v) > 1000 and f.min() == 0 and f.max() == len(v) - 1
    e, cnt = edge_counts(f)
    assert np.all(cnt == 2)                      # closed 2-manifold: every edge shared by exactly 2 triangles
    assert len(v) - len(cnt) + len(f) == 2       # Euler characteristic of a sphere
    # consistent orientation: every directed edge appears exactly once
    assert len(np.unique(e, axis=0)) == len(e)
    # vertices lie on the sphere (linear interpolation error ~ h^2)
    h = ax[1] - ax[0]
    p = v * h - 1.0
    r = np.linalg.norm(p - np.array([0.013, -0.021, 0.007]), axis=1)
    assert np.abs(r - 0.5131).max() < 0.5 * h * h / 0.5131 + 1e-6
    # no duplicated vertices (sharing works)
    assert len(np.unique(np.round(v, 9), axis=0)) == len(v)


def test_vertex_order_is_first_creation_order():
    """Vertex ids increase with the visiting order of the cells that create them (x outer, z inner)."""
    vol, _ = sphere_volume(20)
    v, f = mcubes_ref.marching_cubes(vol, 0.0)
    first_use = np.full(len(v), np.iinfo(np.int64).max)
    np.minimum.at(first_use, f.reshape(-1), np.repeat(np.arange(len(f)), 3))
    assert np.all(first_use < np.iinfo(np.int64).max)
    # the creating cell of vertex v is the cell of the first face using it or an earlier one -> monotone
    assert np.all(np.diff(np.maximum.accumulate(first_use)) >= 0)


def test_empty_and_full_volumes():
    v, f = mcubes_ref.marching_cubes(np.ones((5, 6, 7), dtype=np.float32), 0.0)
    assert v.shape == (0, 3) and f.shape == (0, 3)
    v, f = mcubes_ref.marching_cubes(-np.ones((5, 6, 7), dtype=np.float32), 0.0)
    assert v.shape == (0, 3) and f.shape == (0, 3)


def test_iso_equal_values_count_as_inside():
    """`v <= iso` is inside (the choice recorded in SURVEY.md section 8c) and f1 == f2 gives the midpoint."""
    vol = np.ones((2, 2, 2), dtype=np.float32)
    vol[0, 0, 0] = 0.0
    _, _, cases = mcubes_ref.marching_cubes(vol, 0.0, return_cases=True)
    assert cases[0, 0, 0] == 1


def test_canonical_faces_is_blind_to_numbering_and_low_face_duplicates():
    """The two plausible upstream conventions for lattice edges on the low faces of the volume (one vertex per edge
    vs. a fresh vertex per touching boundary cell) differ in vertex count and ids, never in the surface: the
    canonical face set (faces keyed by the position of the owning lattice edge) is identical."""
    rng = np.random.default_rng(3)
    for vol in (sphere_volume(24, radius=1.2)[0],                       # surface cuts every face of the block
                rng.standard_normal((9, 11, 7)).astype(np.float32)):
        v0, f0 = mcubes_ref.marching_cubes(vol, 0.0)
        v1, f1 = mcubes_ref.marching_cubes(vol, 0.0, dup_low_faces=True)
        assert len(f0) == len(f1) and len(v1) > len(v0)                  # duplicates exist only in the second mode
        assert len(np.unique(v0, axis=0)) == len(v0)                     # mode 0: one vertex per crossed edge
        assert len(np.unique(np.round(v1, 7), axis=0)) == len(v0)        # mode 1: the same edges, some twice (+- 1 ulp)
        c0, u0 = mcubes_ref.canonical_faces(v0, f0)
        c1, u1 = mcubes_ref.canonical_faces(v1, f1)
        assert np.array_equal(u0, u1) and np.array_equal(c0, c1)
        # a random renumbering + face shuffle + cyclic rotation of the same mesh is canonically equal as well
        perm = rng.permutation(len(v0))
        inv = np.empty_like(perm)
        inv[perm] = np.arange(len(perm))
        f2 = np.roll(inv[f0], 1, axis=1)[rng.permutation(len(f0))]
        c2, u2 = mcubes_ref.canonical_faces(v0[perm], f2)
        assert np.array_equal(u0, u2) and np.array_equal(c0, c2)
        # ... while flipping one triangle's orientation is detected
        f3 = f0.copy()
        f3[0] = f3[0][::-1]
        assert not np.array_equal(mcubes_ref.canonical_faces(v0, f3)[0], c0)
