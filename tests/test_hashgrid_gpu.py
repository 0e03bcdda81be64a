"""GPU parity: csrc/hashgrid.cu through the C ABI vs oracle/hashgrid_ref.py.

Tolerances (BASELINE.json north_star: encodings <= 1e-3 relative in fp32): the forward is the same
fp32 formula with a different FMA grouping -> rtol 1e-5 / atol 1e-6 x table scale; the backward is an
atomically accumulated fp32 sum compared with a float64 sum -> 1e-4 relative to the gradient scale.
"""
import numpy as np
import pytest
import torch

from oracle import hashgrid_ref as hr

pytestmark = pytest.mark.gpu

DEFAULT = (16, 8, 22, 32, float(np.exp((np.log(2048) - np.log(32)) / 15)))
SIX_TAPS = lambda e: ((0, 0, 0), (e, 0, 0), (-e, 0, 0), (0, e, 0), (0, -e, 0), (0, 0, e), (0, 0, -e))


def _unit_ball_points(n, seed, lo=0.25, hi=0.75):
    g = torch.Generator().manual_seed(seed)
    p = torch.randn(n, 3, generator=g)
    p = p / p.norm(dim=-1, keepdim=True) * torch.rand(n, 1, generator=g) ** (1 / 3)
    return (p * 0.5 + 0.5) * (hi - lo) + lo


@pytest.fixture(scope="module")
def default_grid(na_lib, cuda_device):
    from neuralangelo_b200 import hashgrid
    tbl = hr.level_table(*DEFAULT)
    g = torch.Generator().manual_seed(11)
    table = torch.rand(tbl["total_entries"] * 8, generator=g) * 2 - 1  # U(-1,1): parity is not vacuous
    desc = hashgrid.make_desc(*DEFAULT)
    return tbl, table, desc, table.to(cuda_device)


@pytest.mark.parametrize("cfg", [(4, 2, 8, 3, 2.0), (5, 4, 10, 4, 1.6), (6, 8, 12, 4, 1.5)])
def test_forward_small_configs(na_lib, cuda_device, cfg):
    from neuralangelo_b200 import hashgrid, _lib
    tbl = hr.level_table(*cfg)
    g = torch.Generator().manual_seed(2)
    table = torch.randn(tbl["total_entries"] * cfg[1], generator=g)
    x = torch.rand(1000, 3, generator=g) * 1.5 - 0.25  # also outside [0,1]: wrap-around indexing
    ref = hr.encode(tbl, table, x)
    desc = hashgrid.make_desc(*cfg)
    out = hashgrid.encode_taps(desc, table.to(cuda_device), x.to(cuda_device), _lib.UnitMap(0.0, 1.0), cfg[0],
                               ((0, 0, 0),))[0]
    torch.testing.assert_close(out.cpu(), ref, rtol=1e-5, atol=1e-5)


def test_forward_default_config(default_grid, cuda_device):
    from neuralangelo_b200 import hashgrid, _lib
    tbl, table, desc, table_d = default_grid
    x = _unit_ball_points(8192, 0)
    ref = hr.encode(tbl, table, x)
    out = hashgrid.encode_taps(desc, table_d, x.to(cuda_device), _lib.UnitMap(0.0, 1.0), 16, ((0, 0, 0),))[0]
    torch.testing.assert_close(out.cpu(), ref, rtol=1e-5, atol=2e-6)


def test_forward_unaligned_output_and_mask(default_grid, cuda_device):
    """[N,131] rows (xyz ++ features, modules.py:94): features start 12 bytes into a 524-byte row."""
    import ctypes
    from neuralangelo_b200 import _lib
    tbl, table, desc, table_d = default_grid
    x = _unit_ball_points(777, 1)
    ref = hr.encode(tbl, table, x, active_levels=5)
    buf = torch.full((777, 131), 7.0, device=cuda_device)
    xd = x.to(cuda_device)
    rc = _lib.lib().na_hashgrid_fwd(ctypes.byref(desc), _lib.ptr(table_d), _lib.ptr(xd), 777, _lib.UnitMap(0.0, 1.0), 5,
                                    ctypes.c_void_p(buf.data_ptr() + 12), 131, _lib.current_stream())
    _lib.check(rc, "na_hashgrid_fwd")
    torch.testing.assert_close(buf[:, 3:].cpu(), ref, rtol=1e-5, atol=2e-6)
    assert torch.all(buf[:, :3] == 7.0)
    assert torch.all(buf[:, 3 + 40:] == 0.0)


@pytest.mark.parametrize("eps", [1.0 / 2049, 1.0 / 33])
def test_taps_forward(default_grid, cuda_device, eps):
    """6-tap offsets in world coordinates with the [-2,2] -> [0,1] map (modules.py:81-82,138-143).
    eps = 1/33 (training start, modules.py:102-107) makes most fine-level taps cross cells."""
    from neuralangelo_b200 import hashgrid, _lib
    tbl, table, desc, table_d = default_grid
    pts = (_unit_ball_points(3000, 4) - 0.5) * 4.0  # world coords, unit ball radius 1
    e = float(np.float32(eps))
    taps = SIX_TAPS(e)
    ref = hr.encode_taps(tbl, table, pts, -2.0, 2.0, taps)
    out = hashgrid.encode_taps(desc, table_d, pts.to(cuda_device), _lib.UnitMap(-2.0, 0.25), 16, taps)
    torch.testing.assert_close(out.cpu(), ref, rtol=1e-5, atol=2e-6)


@pytest.mark.parametrize("n_taps", [1, 7])
def test_backward(default_grid, cuda_device, n_taps):
    from neuralangelo_b200 import hashgrid, _lib
    tbl, table, desc, _ = default_grid
    pts = (_unit_ball_points(5000, 6) - 0.5) * 4.0
    e = float(np.float32(1.0 / 513))
    taps = SIX_TAPS(e)[:n_taps]
    g = torch.Generator().manual_seed(1)
    dout = torch.randn(n_taps, 5000, 128, generator=g)
    ref = torch.zeros(tbl["total_entries"] * 8, dtype=torch.float64)
    for t, off in enumerate(taps):
        p = pts + torch.tensor(off, dtype=torch.float32)
        ref += hr.encode_backward(tbl, hr.unit_map(p, -2.0, 2.0), dout[t], active_levels=12)
    dtable = torch.zeros(tbl["total_entries"] * 8, device=cuda_device)
    hashgrid.scatter_taps(desc, pts.to(cuda_device), _lib.UnitMap(-2.0, 0.25), 12, taps, dout.to(cuda_device), dtable)
    got = dtable.cpu().double()
    scale = ref.abs().max().item()
    assert (got - ref).abs().max().item() <= 1e-4 * scale
    lv12 = tbl["levels"][12]["offset"] * 8
    assert torch.all(got[lv12:] == 0)  # inactive levels receive nothing


def test_autograd_module_matches_oracle(na_lib, cuda_device):
    """tcnn-compatible module: forward values and params.grad through torch autograd."""
    from neuralangelo_b200 import hashgrid
    cfg = dict(otype="HashGrid", n_levels=6, n_features_per_level=4, log2_hashmap_size=12, base_resolution=4,
               per_level_scale=1.5)
    enc = hashgrid.Encoding(3, cfg).to(cuda_device)
    with torch.no_grad():
        enc.params.copy_(torch.randn(enc.params.shape, generator=torch.Generator().manual_seed(0)))
    tbl = hr.level_table(6, 4, 12, 4, 1.5)
    x = torch.rand(513, 3, generator=torch.Generator().manual_seed(9))
    w = torch.randn(513, 24, generator=torch.Generator().manual_seed(10))
    out = enc(x.to(cuda_device))
    (out * w.to(cuda_device)).sum().backward()
    ref = hr.encode(tbl, enc.params.detach().cpu(), x)
    torch.testing.assert_close(out.detach().cpu(), ref, rtol=1e-5, atol=1e-5)
    gref = hr.encode_backward(tbl, x, w)
    assert (enc.params.grad.cpu().double() - gref).abs().max().item() <= 1e-4 * gref.abs().max().item()


def test_full_size_properties(default_grid, cuda_device):
    """BASELINE.json config #2 size (2^22 points): size-independent properties instead of the oracle.
    (i) linearity in the table, (ii) <fwd(T), D> == <T, bwd(D)> (adjointness), (iii) taps == separate calls."""
    from neuralangelo_b200 import hashgrid, _lib
    tbl, table, desc, table_d = default_grid
    n = 1 << 22
    x = _unit_ball_points(n, 0).to(cuda_device)
    um = _lib.UnitMap(0.0, 1.0)
    t2 = torch.rand(table_d.shape, device=cuda_device, generator=torch.Generator(device=cuda_device).manual_seed(3))
    a = hashgrid.encode_taps(desc, table_d, x, um, 16, ((0, 0, 0),))[0]
    b = hashgrid.encode_taps(desc, t2, x, um, 16, ((0, 0, 0),))[0]
    ab = hashgrid.encode_taps(desc, table_d + t2, x, um, 16, ((0, 0, 0),))[0]
    assert (ab - (a + b)).abs().max().item() < 1e-5
    del b, ab, t2
    dout = torch.randn(1, n, 128, device=cuda_device, generator=torch.Generator(device=cuda_device).manual_seed(1))
    dtable = torch.zeros_like(table_d)
    hashgrid.scatter_taps(desc, x, um, 16, ((0, 0, 0),), dout, dtable)
    lhs = (a.double() * dout[0].double()).sum().item()
    rhs = (dtable.double() * table_d.double()).sum().item()
    assert abs(lhs - rhs) <= 1e-5 * abs(lhs) + 1e-3
    del dout, dtable
    e = float(np.float32(1.0 / 8192))
    m = 1 << 20
    taps = hashgrid.encode_taps(desc, table_d, x[:m], um, 16, SIX_TAPS(e))
    for t, off in enumerate(SIX_TAPS(e)):
        xo = x[:m] + torch.tensor(off, device=cuda_device)
        sep = hashgrid.encode_taps(desc, table_d, xo, um, 16, ((0, 0, 0),))[0]
        assert torch.equal(sep, taps[t]) or (sep - taps[t]).abs().max().item() < 1e-6
    del taps, sep
    # (iv) the large-batch 7-evaluation backward (>= 2^21 points: dense levels interleaved across CTAs, the four finest
    # levels scattered in two row-range passes) equals the sum of two half-batch calls, which take the plain path
    n2 = 1 << 21
    g7 = torch.Generator(device=cuda_device).manual_seed(7)
    dout7 = torch.randn(7, 16, n2, 8, device=cuda_device, generator=g7)
    big = torch.zeros_like(table_d)
    hashgrid.scatter_taps_planes(desc, x[:n2], um, 16, SIX_TAPS(e), dout7, big)
    assert hashgrid.bwd_launch_count(desc, n2, 7, 16) == 10 and hashgrid.bwd_launch_count(desc, n2 // 2, 7, 16) == 1
    halves = torch.zeros_like(table_d)
    for h in range(2):
        sl = slice(h * n2 // 2, (h + 1) * n2 // 2)
        hashgrid.scatter_taps_planes(desc, x[:n2][sl], um, 16, SIX_TAPS(e), dout7[:, :, sl].contiguous(), halves)
    scale = halves.abs().max().item()
    assert scale > 0 and (big - halves).abs().max().item() <= 2e-5 * scale     # same terms, another order of fp32 additions
    for l in (0, 5, 6, 11, 12, 15):      # every kind of level: interleaved dense, whole hashed, split hashed
        lo, hi = desc.offset[l] * 8, (desc.offset[l] + desc.entries[l]) * 8
        ref_l2 = halves[lo:hi].double().norm().item()
        assert abs(big[lo:hi].double().norm().item() - ref_l2) <= 1e-6 * ref_l2, l


def test_errors_are_reported(na_lib, cuda_device):
    import ctypes
    from neuralangelo_b200 import hashgrid, _lib
    desc = hashgrid.make_desc(4, 2, 8, 3, 2.0)
    x = torch.rand(4, 3, device=cuda_device)
    rc = _lib.lib().na_hashgrid_fwd(ctypes.byref(desc), None, _lib.ptr(x), 4, _lib.UnitMap(0.0, 1.0), 4, _lib.ptr(x), 8,
                                    _lib.current_stream())
    assert rc == -1 and b"null" in _lib.lib().na_last_error()
    rc = _lib.lib().na_hashgrid_fwd(ctypes.byref(desc), _lib.ptr(x), _lib.ptr(x), 4, _lib.UnitMap(0.0, 1.0), 9, _lib.ptr(x),
                                    8, _lib.current_stream())
    assert rc == -1 and b"active_levels" in _lib.lib().na_last_error()
    with pytest.raises(_lib.NaError):
        hashgrid.encode_taps(desc, torch.zeros(10), torch.rand(4, 3), _lib.UnitMap(0.0, 1.0), 4, ((0, 0, 0),))


def test_plane_layout_matches_row_layout(default_grid, cuda_device):
    """[taps][levels][points][F] planes hold the same numbers as the [taps][points][L*F] rows, forward and backward."""
    from neuralangelo_b200 import hashgrid, _lib
    tbl, table, desc, table_d = default_grid
    pts = ((_unit_ball_points(4099, 8) - 0.5) * 4.0).to(cuda_device)
    um = _lib.UnitMap(-2.0, 0.25)
    taps = SIX_TAPS(float(np.float32(1.0 / 2049)))
    rows = hashgrid.encode_taps(desc, table_d, pts, um, 13, taps)                       # [7, N, 128]
    planes = hashgrid.encode_taps_planes(desc, table_d, pts, um, 13, taps)              # [7, 16, N, 8]
    assert torch.equal(planes.permute(0, 2, 1, 3).reshape(7, -1, 128), rows)
    g = torch.Generator(device=cuda_device).manual_seed(2)
    dout = torch.randn(7, 4099, 128, device=cuda_device, generator=g)
    d_rows = torch.zeros_like(table_d)
    hashgrid.scatter_taps(desc, pts, um, 13, taps, dout, d_rows)
    d_planes = torch.zeros_like(table_d)
    hashgrid.scatter_taps_planes(desc, pts, um, 13, taps, dout.view(7, -1, 16, 8).permute(0, 2, 1, 3).contiguous(), d_planes)
    assert (d_rows - d_planes).abs().max().item() <= 1e-5 * d_rows.abs().max().item()
