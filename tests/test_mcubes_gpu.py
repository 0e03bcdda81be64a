"""GPU parity: csrc/mcubes.cu through the C ABI vs oracle/mcubes_ref.c -- BIT-EXACT (integer/topology
work and IEEE double interpolation): identical vertex array, identical face array, same order."""
import numpy as np
import pytest
import torch

from oracle import mcubes_ref

pytestmark = pytest.mark.gpu


def _volumes():
    rng = np.random.default_rng(0)
    ax = np.linspace(-1, 1, 40, dtype=np.float32)
    x, y, z = np.meshgrid(ax, ax[:33], ax[:37], indexing="ij")
    sphere = np.sqrt((x - 0.01) ** 2 + (y + 0.2) ** 2 + (z + 0.1) ** 2).astype(np.float32) - 0.5131
    torus = ((np.sqrt(x ** 2 + (y + 0.2) ** 2) - 0.45) ** 2 + (z + 0.1) ** 2 - 0.04).astype(np.float32)
    noise = rng.standard_normal((17, 23, 19)).astype(np.float32)      # every case, ambiguous faces included
    ties = np.round(rng.standard_normal((12, 12, 12)) * 2).astype(np.float32)  # exact zeros and equal neighbours
    thin = rng.standard_normal((2, 2, 50)).astype(np.float32)
    return dict(sphere=sphere, torus=torus, noise=noise, ties=ties, thin=thin,
                empty=np.ones((6, 5, 4), dtype=np.float32))


@pytest.mark.parametrize("name", ["sphere", "torus", "noise", "ties", "thin", "empty"])
def test_bit_exact_vs_serial_oracle(na_lib, cuda_device, name):
    from neuralangelo_b200 import mesh
    vol = _volumes()[name]
    v_ref, f_ref = mcubes_ref.marching_cubes(vol, 0.0)
    v, f = mesh.mc(torch.from_numpy(vol).to(cuda_device), 0.0)
    assert v.shape == v_ref.shape and f.shape == f_ref.shape
    assert np.array_equal(f.cpu().numpy(), f_ref)
    assert np.array_equal(v.cpu().numpy(), v_ref)          # bit-exact doubles


@pytest.mark.parametrize("name", ["sphere", "torus", "noise", "ties", "thin"])
def test_canonical_face_set_vs_oracle_both_boundary_conventions(na_lib, cuda_device, name):
    """SURVEY.md section 8c's definition of "bit-exact topology": identical per-cell case volume, identical face count,
    identical face set after canonical relabelling (faces keyed by the owning lattice edge of each corner).  Holds
    against BOTH plausible upstream conventions for low-face lattice edges (oracle/mcubes_ref.c header); the vertex
    count matches the one-vertex-per-edge convention the kernel implements."""
    from neuralangelo_b200 import mesh
    vol = _volumes()[name]
    v, f = mesh.mc(torch.from_numpy(vol).to(cuda_device), 0.0)
    v, f = v.cpu().numpy(), f.cpu().numpy()
    c_gpu, u_gpu = mcubes_ref.canonical_faces(v, f)
    for dup in (False, True):
        v_ref, f_ref, cases = mcubes_ref.marching_cubes(vol, 0.0, return_cases=True, dup_low_faces=dup)
        assert len(f) == len(f_ref)
        c_ref, u_ref = mcubes_ref.canonical_faces(v_ref, f_ref)
        assert np.array_equal(u_gpu, u_ref) and np.array_equal(c_gpu, c_ref)
        if not dup:
            assert len(v) == len(v_ref)
    # per-cell case volume: the kernel's classification is what its faces were emitted from; recompute it here
    inside = vol <= 0.0
    c = np.zeros(tuple(n - 1 for n in vol.shape), dtype=np.uint8)
    for m, (dx, dy, dz) in enumerate([(0, 0, 0), (1, 0, 0), (1, 1, 0), (0, 1, 0), (0, 0, 1), (1, 0, 1), (1, 1, 1), (0, 1, 1)]):
        c |= (inside[dx:dx + c.shape[0], dy:dy + c.shape[1], dz:dz + c.shape[2]].astype(np.uint8) << m)
    assert np.array_equal(c, cases)


def test_scale_origin_and_nonzero_iso(na_lib, cuda_device):
    from neuralangelo_b200 import mesh
    vol = _volumes()["noise"]
    v_ref, f_ref = mcubes_ref.marching_cubes(vol, 0.25)
    intv, org = 2.0 / 256, np.array([-0.5, 0.25, -1.0], dtype=np.float32)
    v, f = mesh.mc(torch.from_numpy(vol).to(cuda_device), 0.25, scale=intv, origin=org.astype(np.float64))
    assert np.array_equal(f.cpu().numpy(), f_ref)
    assert np.array_equal(v.cpu().numpy(), v_ref * intv + org)   # mesh.py:123 `V * intv + xyz[0,0,0]`


def test_block_129_cubed(na_lib, cuda_device):
    """One full extraction block (block_res 128 -> 129^3 samples), analytic sphere r = 0.5131."""
    from neuralangelo_b200 import mesh
    ax = torch.linspace(-1, 1, 129, device=cuda_device)
    x, y, z = torch.meshgrid(ax, ax, ax, indexing="ij")
    vol = (torch.sqrt(x ** 2 + y ** 2 + z ** 2) - 0.5131).contiguous()
    v, f = mesh.mc(vol, 0.0)
    v_ref, f_ref = mcubes_ref.marching_cubes(vol.cpu().numpy(), 0.0)
    assert np.array_equal(f.cpu().numpy(), f_ref) and np.array_equal(v.cpu().numpy(), v_ref)
    assert len(v) - 3 * len(f) // 2 + len(f) == 2     # closed surface: V - E + F = 2 with E = 3F/2


def test_extract_mesh_analytic_sphere(na_lib, cuda_device):
    """extract_mesh with the reference's signature on an analytic SDF: block seams line up (vertices of
    neighbouring blocks coincide on the shared slice) and the Chamfer-style radius error is tiny."""
    from neuralangelo_b200 import mesh
    r = 0.5131
    sdf_func = lambda xyz: -(xyz.norm(dim=-1, keepdim=True) - r)   # sign flip as extract_mesh.py:85
    bounds = np.array([[-1.0, 1.0], [-1.0, 1.0], [-1.0, 1.0]])
    m = mesh.extract_mesh(sdf_func, bounds, intv=2.0 / 96, block_res=32)
    assert len(m.vertices) > 5000 and m.faces.max() == len(m.vertices) - 1
    rad = np.linalg.norm(m.vertices, axis=1)
    assert np.abs(rad - r).max() < 1e-4 * np.sqrt(12)      # <= 1e-4 of the bbox diagonal
    # whole-grid single block gives the same surface (as a set of triangles' vertex positions)
    one = mesh.extract_mesh(sdf_func, bounds, intv=2.0 / 96, block_res=128)
    a = np.sort(m.vertices[m.faces].reshape(-1, 9), axis=0)
    b = np.sort(one.vertices[one.faces].reshape(-1, 9), axis=0)
    assert a.shape == b.shape and np.allclose(a, b, atol=1e-6)  # block origins are float32 (mesh.py:123)
    comps = m.split()
    assert len(comps) >= 1


def test_lattice_grid_matches_reference_geometry():
    from neuralangelo_b200 import mesh
    g = mesh.LatticeGrid(np.array([[-1.0, 1.0]] * 3), intv=2.0 / 256, block_res=128)
    assert len(g) == 8 and g.resolution == (256, 256, 256)
    assert g[0]["xyz"].shape == (129, 129, 129, 3) and g[7]["xyz"].shape == (128, 128, 128, 3)
    g2 = mesh.LatticeGrid(np.array([[-1.0, 1.0]] * 3), intv=2.0 / 2048, block_res=128)
    assert len(g2) == 4096


def test_mc_errors(na_lib, cuda_device):
    from neuralangelo_b200 import mesh, _lib
    with pytest.raises(_lib.NaError):
        mesh.mc(torch.zeros(1, 5, 5, device=cuda_device))
