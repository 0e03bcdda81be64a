"""CPU restatement of tiny-cuda-nn's multi-resolution hash grid -- TEST INFRASTRUCTURE ONLY.

PARITY UNPINNED: the arithmetic lives in a third-party dependency that is NOT vendored under the
reference tree: tiny-cuda-nn torch bindings, `git+https://github.com/NVlabs/tiny-cuda-nn/
#subdirectory=bindings/torch`, pinned version: none (floating HEAD; reference requirements.txt:3).
Reference call sites: projects/neuralangelo/utils/modules.py:16 (import), :42-50 (config +
construction), :85 (forward); projects/nerf/models/ingp.py:56,71.  The reference holds no test,
fixture or golden vector at this boundary and tcnn cannot be installed here (no network, no GPU),
so this file restates tcnn's published algorithm (grid.h: grid_scale, grid_resolution, grid_index,
pos_fract, kernel_grid, kernel_grid_backward) and is anchored only on hand-derived known answers
(tests/test_oracle_hashgrid.py) -- not on outputs of tcnn itself.

Spec followed:
  1. log2s = log2f(float(per_level_scale)); scale_l = exp2f(l*log2s)*base - 1 (float32);
     res_l = ceil(scale_l) + 1.
  2. entries_l = min(round_up_8(res_l^3, saturating at 2^31-1), 2^log2_hashmap_size); offsets = running sum.
  3. pos = fma(scale, x, 0.5) in float32; cell = floor(pos); w = pos - cell; no clamping.
  4. corner index: dense stride walk (x + y*res + z*res^2) or, when the level is hashed
     (entries < res^3), x ^ y*2654435761 ^ z*805459861 in uint32; then modulo entries_l.
  5. out[n, l*F+f] = sum over 8 corners of weight * table[(offset_l+idx)*F+f]   (fp32 here).
  6. backward: dtable[(offset_l+idx)*F+f] += weight * dout[n, l*F+f].
"""
import numpy as np
import torch

PRIME_Y = 2654435761
PRIME_Z = 805459861
U32 = 0xFFFFFFFF


def level_table(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    """Item 1-2 of the spec.  Transcendentals in float64, rounded once to float32 (= a correctly
    rounded log2f/exp2f; tcnn's device exp2f may differ by an ulp -- unverifiable here)."""
    pls = np.float32(per_level_scale)
    log2s = np.float32(np.log2(np.float64(pls)))
    hashmap = 1 << log2_hashmap_size
    levels = []
    offset = 0
    for l in range(n_levels):
        e = np.float32(np.exp2(np.float64(np.float32(l) * log2s)))
        scale = np.float32(e * np.float32(base_resolution) - np.float32(1.0))
        res = int(np.ceil(scale)) + 1
        dense = res ** 3
        params = min(dense, U32 // 2)
        params = (params + 7) // 8 * 8
        entries = min(params, hashmap)
        levels.append(dict(scale=float(scale), res=res, entries=entries, offset=offset, hashed=dense > entries))
        offset += entries
    return dict(levels=levels, n_levels=n_levels, n_features=n_features, total_entries=offset)


def _corner_indices(lv, cell):
    """cell: int64 [N,3] holding uint32 values; returns (idx[8,N] int64, bits[8,3])."""
    res, entries = lv["res"], lv["entries"]
    out = []
    for k in range(8):
        cx = (cell[:, 0] + (k & 1)) & U32
        cy = (cell[:, 1] + ((k >> 1) & 1)) & U32
        cz = (cell[:, 2] + ((k >> 2) & 1)) & U32
        if lv["hashed"]:
            idx = cx ^ ((cy * PRIME_Y) & U32) ^ ((cz * PRIME_Z) & U32)
        else:
            idx = (cx + ((cy * res) & U32) + ((cz * ((res * res) & U32)) & U32)) & U32
        out.append(idx % entries)
    return torch.stack(out, 0)


def _locate(lv, x):
    """x float32 [N,3] -> (cell int64 [N,3] as uint32 values, w float32 [N,3])."""
    scale = lv["scale"]
    # fma(scale, x, 0.5): the product of two float32 is exact in float64, adding 0.5 stays exact for
    # |pos| < 2^29 or so, hence a single rounding to float32 -- the fused-multiply-add result.
    pos = (x.double() * float(np.float32(scale)) + 0.5).float()
    fl = torch.floor(pos)
    w = pos - fl
    cell = fl.to(torch.int64) & U32  # (uint32)(int)floor
    return cell, w


def _weights(w):
    ws = []
    for k in range(8):
        wx = w[:, 0] if (k & 1) else 1.0 - w[:, 0]
        wy = w[:, 1] if (k & 2) else 1.0 - w[:, 1]
        wz = w[:, 2] if (k & 4) else 1.0 - w[:, 2]
        ws.append(wx * wy * wz)
    return torch.stack(ws, 0)  # [8,N]


def encode(tbl, table, x, active_levels=None, chunk=1 << 18):
    """table: float32 [total_entries*F]; x float32 [N,3] (unit cube coords) -> float32 [N, L*F]."""
    L, F = tbl["n_levels"], tbl["n_features"]
    active = L if active_levels is None else active_levels
    table = table.detach().reshape(-1, F)
    x = x.detach().float().reshape(-1, 3)
    out = torch.zeros(x.shape[0], L * F, dtype=torch.float32)
    for s in range(0, x.shape[0], chunk):
        xs = x[s:s + chunk]
        for l in range(active):
            lv = tbl["levels"][l]
            cell, w = _locate(lv, xs)
            idx = _corner_indices(lv, cell) + lv["offset"]
            ws = _weights(w)
            acc = torch.zeros(xs.shape[0], F, dtype=torch.float32)
            for k in range(8):
                acc += ws[k][:, None] * table[idx[k]]
            out[s:s + chunk, l * F:(l + 1) * F] = acc
    return out


def encode_backward(tbl, x, dout, active_levels=None, chunk=1 << 18, dtype=torch.float64):
    """Returns dtable [total_entries*F] accumulated in `dtype` (float64 by default: the atomically
    accumulated fp32 GPU result is compared against this within a summation-order tolerance)."""
    L, F = tbl["n_levels"], tbl["n_features"]
    active = L if active_levels is None else active_levels
    x = x.detach().float().reshape(-1, 3)
    dout = dout.detach().float().reshape(-1, L * F)
    dtable = torch.zeros(tbl["total_entries"], F, dtype=dtype)
    for s in range(0, x.shape[0], chunk):
        xs = x[s:s + chunk]
        for l in range(active):
            lv = tbl["levels"][l]
            cell, w = _locate(lv, xs)
            idx = _corner_indices(lv, cell) + lv["offset"]
            ws = _weights(w)
            d = dout[s:s + chunk, l * F:(l + 1) * F]
            for k in range(8):
                dtable.index_add_(0, idx[k], (ws[k][:, None] * d).to(dtype))
    return dtable.reshape(-1)


def unit_map(points, vol_min, vol_max):
    """modules.py:81-82 normalisation, as the product does it: (p - min) * (1/(max-min)) in float32."""
    inv = np.float32(1.0) / np.float32(vol_max - vol_min)
    return (points.float() - float(np.float32(vol_min))) * float(inv)


def encode_taps(tbl, table, points, vol_min, vol_max, tap_offsets, active_levels=None):
    """[n_taps, N, L*F]: encode(unit_map(points + tap)) with the add done in float32 first
    (modules.py:138-143 `self.sdf(x + eps_x)`, :164-167 `x + k1 * eps`)."""
    outs = []
    for off in tap_offsets:
        p = points.float().reshape(-1, 3) + torch.tensor(off, dtype=torch.float32)
        outs.append(encode(tbl, table, unit_map(p, vol_min, vol_max), active_levels))
    return torch.stack(outs, 0)


def init_table(tbl, seed=1337, amplitude=1e-4):
    gen = torch.Generator(device="cpu").manual_seed(seed)
    n = tbl["total_entries"] * tbl["n_features"]
    return (torch.rand(n, generator=gen, dtype=torch.float32) * 2.0 - 1.0) * amplitude
