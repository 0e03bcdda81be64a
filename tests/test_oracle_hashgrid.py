"""CPU: pins oracle/hashgrid_ref.py against an independent scalar restatement and hand-derived
known answers, and the library's host-side level table against the oracle's (no GPU needed)."""
import math

import numpy as np
import pytest
import torch

from oracle import hashgrid_ref as hr

U32 = 0xFFFFFFFF


def scalar_encode(tbl, table, x):
    """Scalar python restatement (independent of the vectorised oracle): one point, loops only."""
    F = tbl["n_features"]
    out = []
    for lv in tbl["levels"]:
        s = np.float32(lv["scale"])
        pos = [np.float32(np.float64(s) * np.float64(np.float32(c)) + 0.5) for c in x]
        cell = [int(math.floor(p)) & U32 for p in pos]
        w = [np.float32(p - np.float32(math.floor(p))) for p in pos]
        acc = np.zeros(F, dtype=np.float32)
        for k in range(8):
            c = [(cell[d] + ((k >> d) & 1)) & U32 for d in range(3)]
            if lv["hashed"]:
                idx = c[0] ^ ((c[1] * 2654435761) & U32) ^ ((c[2] * 805459861) & U32)
            else:
                idx = (c[0] + c[1] * lv["res"] + c[2] * lv["res"] * lv["res"]) & U32
            idx %= lv["entries"]
            wk = np.float32(1.0)
            for d in range(3):
                wk = np.float32(wk * (w[d] if (k >> d) & 1 else np.float32(1.0) - w[d]))
            acc += wk * table[(lv["offset"] + idx) * F:(lv["offset"] + idx + 1) * F]
        out.append(acc)
    return np.concatenate(out)


def test_level_table_default_config():
    """SURVEY.md section 8 geometry table for 16 levels, dict_size 22, dim 8, 32 -> 2048."""
    g = np.exp((np.log(2048) - np.log(32)) / 15)  # modules.py:38-41
    t = hr.level_table(16, 8, 22, 32, g)
    assert [lv["res"] for lv in t["levels"]] == [32, 43, 56, 74, 98, 129, 169, 223, 295, 389, 513, 676, 892, 1177,
                                                 1553, 2049]
    assert [lv["entries"] for lv in t["levels"]][:7] == [32768, 79512, 175616, 405224, 941192, 2146696, 4194304]
    assert t["total_entries"] == 45724048
    assert [lv["hashed"] for lv in t["levels"]] == [False] * 6 + [True] * 10


def test_library_level_table_matches_oracle(na_lib):
    from neuralangelo_b200 import hashgrid
    for cfg in [(16, 8, 22, 32, np.exp((np.log(2048) - np.log(32)) / 15)), (16, 4, 20, 32, 1.3195079107728942),
                (8, 2, 14, 16, 2.0), (4, 4, 19, 16, 1.5), (12, 8, 21, 32, 1.4)]:
        t = hr.level_table(*cfg)
        d = hashgrid.desc_to_dict(hashgrid.make_desc(*cfg))
        assert d["total_entries"] == t["total_entries"]
        for l, lv in enumerate(t["levels"]):
            assert np.float32(lv["scale"]) == np.float32(d["scale"][l])
            assert (lv["res"], lv["entries"], lv["offset"], int(lv["hashed"])) == \
                (d["resolution"][l], d["entries"][l], d["offset"][l], d["hashed"][l])


def test_known_answer_tiny_grid():
    """Hand-derived: 1 dense level, scale 1, res 2, 8 entries, F=2; table[e] = (e, 10e)."""
    t = hr.level_table(1, 2, 4, 2, 2.0)
    lv = t["levels"][0]
    assert (lv["scale"], lv["res"], lv["entries"], lv["hashed"]) == (1.0, 2, 8, False)
    table = torch.tensor([[e, 10.0 * e] for e in range(8)], dtype=torch.float32).reshape(-1)
    # x = (0.25, 0.5, 0.75): pos = (0.75, 1.0, 1.25) -> cell (0,1,1), w = (0.75, 0, 0.25)
    # corners with non-zero weight: (0,1,1)->idx 6 w .25*.75 ; (1,1,1)->7 w .75*.75 ;
    #   (0,1,2)-> 0+2+8=10 -> 10%8=2 w .25*.25 ; (1,1,2)-> 11%8=3 w .75*.25
    expect = 6 * 0.1875 + 7 * 0.5625 + 2 * 0.0625 + 3 * 0.1875
    out = hr.encode(t, table, torch.tensor([[0.25, 0.5, 0.75]]))
    assert out.shape == (1, 2)
    np.testing.assert_allclose(out[0].numpy(), [expect, 10 * expect], rtol=1e-6)


def test_known_answer_hash():
    """Hashed level: the corner (1,2,3) must land on (1 ^ 2*2654435761 ^ 3*805459861) mod 2^4."""
    t = hr.level_table(2, 2, 4, 2, 2.0)
    lv = t["levels"][1]
    assert (lv["scale"], lv["res"], lv["entries"], lv["hashed"]) == (3.0, 4, 16, True)
    h = (1 ^ ((2 * 2654435761) & U32) ^ ((3 * 805459861) & U32)) % 16
    table = torch.zeros(t["total_entries"] * 2)
    table[(lv["offset"] + h) * 2] = 1.0
    # pos = 3x+0.5 = (1.0, 2.0, 3.0) exactly -> cell (1,2,3), weights 0 -> corner 0 has weight 1
    x = torch.tensor([[0.5 / 3, 1.5 / 3, 2.5 / 3]], dtype=torch.float32)
    cell, w = hr._locate(lv, x)
    if cell.tolist() == [[1, 2, 3]]:
        out = hr.encode(t, table, x)
        np.testing.assert_allclose(out[0, 2].item(), float(hr._weights(w)[0, 0]), rtol=1e-6)
    idx = hr._corner_indices(lv, torch.tensor([[1, 2, 3]]))
    assert idx[0, 0].item() == h


@pytest.mark.parametrize("cfg", [(3, 2, 6, 4, 1.7), (4, 4, 8, 3, 2.0), (2, 8, 5, 5, 1.3)])
def test_vectorised_oracle_matches_scalar(cfg):
    t = hr.level_table(*cfg)
    g = torch.Generator().manual_seed(3)
    table = torch.randn(t["total_entries"] * t["n_features"], generator=g)
    x = torch.rand(32, 3, generator=g) * 1.4 - 0.2  # includes points outside [0,1] (wrap-around path)
    out = hr.encode(t, table, x)
    for i in range(x.shape[0]):
        ref = scalar_encode(t, table.numpy(), x[i].numpy())
        np.testing.assert_allclose(out[i].numpy(), ref, rtol=2e-5, atol=2e-6)


def test_backward_is_adjoint_of_forward():
    t = hr.level_table(4, 4, 8, 3, 2.0)
    g = torch.Generator().manual_seed(5)
    table = torch.randn(t["total_entries"] * 4, generator=g)
    x = torch.rand(200, 3, generator=g)
    dout = torch.randn(200, 16, generator=g)
    lhs = (hr.encode(t, table, x).double() * dout.double()).sum()
    rhs = (hr.encode_backward(t, x, dout) * table.double()).sum()
    assert abs(lhs - rhs) < 1e-4 * max(1.0, abs(lhs))


def test_active_levels_mask():
    t = hr.level_table(4, 2, 8, 3, 2.0)
    table = torch.ones(t["total_entries"] * 2)
    out = hr.encode(t, table, torch.rand(5, 3), active_levels=2)
    assert torch.all(out[:, 4:] == 0) and torch.all(out[:, :4] > 0.99)
