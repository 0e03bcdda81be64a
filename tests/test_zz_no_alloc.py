"""GPU test of `na_mc_run_static`, the marching-cubes entry point for callers that cannot allocate
device memory (its scratch is a fixed array inside the library instead of a caller-provided buffer).

Collected last on purpose (file name), so a run that stops at the first failure has already run
everything else.  The check: called through the C ABI with caller-owned input/output tensors, twice in
a row on different volumes (the second smaller than the first, so stale scratch from the first call
would show), its vertices, triangles and counts must equal -- bit for bit, the tolerance the other
marching-cubes tests use -- what the regular na_mc_count/na_mc_emit pair returns.  It must pass both
in a fresh process and after the rest of the suite has used the library.
"""
import ctypes

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run_static(lib, vol, iso, scale, origin, cap_v, cap_f):
    from neuralangelo_b200 import _lib
    dev = vol.device
    verts = torch.full((cap_v, 3), float("nan"), dtype=torch.float64, device=dev)
    faces = torch.full((cap_f, 3), -1, dtype=torch.int32, device=dev)
    counts = torch.full((2,), 123456, dtype=torch.int32, device=dev)
    org = (ctypes.c_double * 3)(*origin)
    nx, ny, nz = vol.shape
    rc = lib.na_mc_run_static(_lib.ptr(vol), nx, ny, nz, iso, scale, org, _lib.ptr(verts), cap_v, _lib.ptr(faces), cap_f,
                              _lib.ptr(counts), _lib.current_stream())
    _lib.check(rc, "na_mc_run_static")
    n_v, n_f = (int(c) for c in counts.cpu())
    return verts, faces, n_v, n_f


def test_static_scratch_matches_regular_entry_points_twice(na_lib, cuda_device):
    from neuralangelo_b200 import mesh
    rng = np.random.default_rng(7)
    big = torch.from_numpy(rng.standard_normal((65, 65, 65)).astype(np.float32)).to(cuda_device)   # the maximum size
    ax = torch.linspace(-1, 1, 40, device=cuda_device)
    x, y, z = torch.meshgrid(ax, ax[:31], ax[:23], indexing="ij")
    small = (torch.sqrt(x ** 2 + y ** 2 + z ** 2) - 0.5131).contiguous()
    assert (big.shape[0] - 1) ** 3 == na_lib.na_mc_static_max_cells()
    for vol, iso, scale, origin in ((big, 0.1, 1.0, (0.0, 0.0, 0.0)), (small, 0.0, 2.0 / 39, (-1.0, -1.0, -1.0)),
                                    (big, 0.1, 1.0, (0.0, 0.0, 0.0))):
        v_ref, f_ref = mesh.mc(vol, iso, scale=scale, origin=origin)          # existing path: caller-owned scratch
        cap_v, cap_f = v_ref.shape[0] + 7, f_ref.shape[0] + 5
        verts, faces, n_v, n_f = _run_static(na_lib, vol, iso, scale, origin, cap_v, cap_f)
        assert (n_v, n_f) == (v_ref.shape[0], f_ref.shape[0])
        assert torch.equal(verts[:n_v], v_ref)                                   # bit-exact float64
        assert torch.equal(faces[:n_f].long(), f_ref)
        assert torch.isnan(verts[n_v:]).all() and (faces[n_f:] == -1).all()      # nothing written past the counts


def test_static_scratch_respects_capacities_and_rejects_large_volumes(na_lib, cuda_device):
    from neuralangelo_b200 import _lib, mesh
    rng = np.random.default_rng(3)
    vol = torch.from_numpy(rng.standard_normal((20, 20, 20)).astype(np.float32)).to(cuda_device)
    v_ref, f_ref = mesh.mc(vol, 0.0)
    verts, faces, n_v, n_f = _run_static(na_lib, vol, 0.0, 1.0, (0.0, 0.0, 0.0), 100, 50)   # too small on purpose
    assert (n_v, n_f) == (v_ref.shape[0], f_ref.shape[0]) and n_v > 100 and n_f > 50        # true counts still reported
    assert torch.equal(verts, v_ref[:100]) and torch.equal(faces.long(), f_ref[:50])
    too_big = torch.zeros(66, 65, 65, device=cuda_device)
    counts = torch.zeros(2, dtype=torch.int32, device=cuda_device)
    rc = na_lib.na_mc_run_static(_lib.ptr(too_big), 66, 65, 65, 0.0, 1.0, None, None, 0, None, 0, _lib.ptr(counts),
                                 _lib.current_stream())
    assert rc == -3
    msg = na_lib.na_last_error().decode()
    assert "exceed" in msg and "262144" in msg
