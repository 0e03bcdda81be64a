"""Multi-resolution hash grid: drop-in for `tinycudann.Encoding(3, {otype: HashGrid, ...})`.

Mirrors the call surface the reference uses (projects/neuralangelo/utils/modules.py:42-50
constructs it, modules.py:85 calls it, the trainer/checkpointer see a flat fp32 parameter
named `params`): `Encoding(n_input_dims, encoding_config, seed=1337, dtype=None)` returns an
`nn.Module` whose `forward(x[N,3] in [0,1]) -> [N, n_levels*n_features]`.  Differences from
tcnn, on purpose: features are produced in fp32 from the fp32 master table (tcnn re-casts the
whole table to fp16 on every call and returns fp16), and inputs are not padded to a batch
granularity.  All arithmetic runs in libna_b200.so (csrc/hashgrid.cu).
"""
import ctypes

import torch

from . import _lib

_F32 = torch.float32


def make_desc(n_levels, n_features, log2_hashmap_size, base_resolution, per_level_scale):
    """Level table (scale/resolution/entries/offset/hashed per level), computed by the library."""
    desc = _lib.HashGridDesc()
    rc = _lib.lib().na_hashgrid_make_desc(int(n_levels), int(n_features), int(log2_hashmap_size),
                                          int(base_resolution), float(per_level_scale), ctypes.byref(desc))
    _lib.check(rc, "na_hashgrid_make_desc")
    return desc


def desc_to_dict(desc):
    n = desc.n_levels
    return dict(n_levels=n, n_features=desc.n_features, total_entries=desc.total_entries,
                scale=[desc.scale[i] for i in range(n)], resolution=[desc.resolution[i] for i in range(n)],
                entries=[desc.entries[i] for i in range(n)], offset=[desc.offset[i] for i in range(n)],
                hashed=[desc.hashed[i] for i in range(n)])


def _taps_array(tap_offsets):
    flat = [float(v) for row in tap_offsets for v in row]
    return (ctypes.c_float * len(flat))(*flat), len(flat) // 3


def _as_points(x):
    if x.dtype != _F32:
        x = x.float()
    return x.contiguous()


def encode_taps(desc, table, x, unit_map, active_levels, tap_offsets, out=None):
    """out[t, n, :] = encode(x[n] + tap_offsets[t]); raw (non-autograd) kernel call."""
    _lib.require_cuda(table, x)
    x = _as_points(x)
    n = x.shape[0]
    width = desc.n_levels * desc.n_features
    taps, n_taps = _taps_array(tap_offsets)
    if out is None:
        out = torch.empty((n_taps, n, width), dtype=_F32, device=x.device)
    rc = _lib.lib().na_hashgrid_taps_fwd(ctypes.byref(desc), _lib.ptr(table), _lib.ptr(x), n, unit_map,
                                         int(active_levels), taps, n_taps, _lib.ptr(out), out.stride(1),
                                         out.stride(0), _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_fwd")
    return out


def scatter_taps(desc, x, unit_map, active_levels, tap_offsets, dout, dtable):
    """dtable += d encode(x + taps) / d table applied to dout[t, n, :] (atomic accumulation)."""
    _lib.require_cuda(dtable, x, dout)
    x = _as_points(x)
    if dout.dtype != _F32:
        dout = dout.float()
    if dout.stride(-1) != 1:
        dout = dout.contiguous()
    taps, n_taps = _taps_array(tap_offsets)
    assert dout.shape[0] == n_taps and dout.shape[1] == x.shape[0]
    rc = _lib.lib().na_hashgrid_taps_bwd(ctypes.byref(desc), _lib.ptr(x), x.shape[0], unit_map, int(active_levels),
                                         taps, n_taps, _lib.ptr(dout), dout.stride(1), dout.stride(0),
                                         _lib.ptr(dtable), _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_bwd")
    return dtable


def encode_taps_planes(desc, table, x, unit_map, active_levels, tap_offsets, out=None):
    """Level-major feature planes: out[t, l, n, f] = feature f of level l at x[n] + tap_offsets[t]."""
    _lib.require_cuda(table, x)
    x = _as_points(x)
    n, L, F = x.shape[0], desc.n_levels, desc.n_features
    taps, n_taps = _taps_array(tap_offsets)
    if out is None:
        out = torch.empty((n_taps, L, n, F), dtype=_F32, device=x.device)
    assert out.is_contiguous() and out.shape == (n_taps, L, n, F)
    rc = _lib.lib().na_hashgrid_taps_fwd_strided(ctypes.byref(desc), _lib.ptr(table), _lib.ptr(x), n, unit_map,
                                                 int(active_levels), taps, n_taps, _lib.ptr(out), F, L * n * F, n * F,
                                                 _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_fwd_strided")
    return out


def bwd_launch_count(desc, n_points, n_taps, active_levels):
    """Kernel launches of one backward call on the axis-factorised path (mirrors launch_bwd in csrc/hashgrid.cu): for a
    large batch one launch for the leading dense levels (interleaved across CTAs), one for the other levels scattered
    whole, two row-range passes for each of the (at most four) finest hashed levels of a 7-evaluation batch whose
    gradient exceeds 96 MiB."""
    split = active_levels
    if n_points >= (1 << 21) and n_taps == 7:
        while split > 0 and split + 4 > active_levels and desc.hashed[split - 1] and \
                desc.entries[split - 1] * desc.n_features * 4 > (96 << 20):
            split -= 1
    dense = 0
    if n_points >= (1 << 21):
        while dense < split and not desc.hashed[dense]:
            dense += 1
    if dense <= 1:
        dense = 0
    whole = (1 if dense else 0) + (1 if split > dense else 0)     # interleaved dense levels, remaining whole levels
    return whole + 2 * (active_levels - split)


def scatter_taps_planes(desc, x, unit_map, active_levels, tap_offsets, dout, dtable):
    """dtable += backward of encode_taps_planes for dout[t, l, n, f] (contiguous)."""
    _lib.require_cuda(dtable, x, dout)
    x = _as_points(x)
    n, L, F = x.shape[0], desc.n_levels, desc.n_features
    taps, n_taps = _taps_array(tap_offsets)
    assert dout.is_contiguous() and dout.shape == (n_taps, L, n, F) and dout.dtype == _F32
    rc = _lib.lib().na_hashgrid_taps_bwd_strided(ctypes.byref(desc), _lib.ptr(x), n, unit_map, int(active_levels), taps,
                                                 n_taps, _lib.ptr(dout), F, L * n * F, n * F, _lib.ptr(dtable),
                                                 _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_bwd_strided")
    return dtable


def encode_taps_strided(desc, table, x, unit_map, active_levels, tap_offsets, out, point_stride, tap_stride, level_stride):
    """out[t * tap_stride + l * level_stride + n * point_stride + f] (float offsets into `out`) = feature f of level l
    at x[n] + tap_offsets[t]; the caller owns the layout (e.g. the chunk planes the SDF MLP reads)."""
    _lib.require_cuda(table, x, out)
    x = _as_points(x)
    taps, n_taps = _taps_array(tap_offsets)
    assert out.dtype == _F32
    rc = _lib.lib().na_hashgrid_taps_fwd_strided(ctypes.byref(desc), _lib.ptr(table), _lib.ptr(x), x.shape[0], unit_map,
                                                 int(active_levels), taps, n_taps, _lib.ptr(out), int(point_stride),
                                                 int(tap_stride), int(level_stride), _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_fwd_strided")
    return out


def scatter_taps_strided(desc, x, unit_map, active_levels, tap_offsets, dout, point_stride, tap_stride, level_stride, dtable):
    """dtable += backward of encode_taps_strided for the same layout of `dout`."""
    _lib.require_cuda(dtable, x, dout)
    x = _as_points(x)
    taps, n_taps = _taps_array(tap_offsets)
    assert dout.dtype == _F32
    rc = _lib.lib().na_hashgrid_taps_bwd_strided(ctypes.byref(desc), _lib.ptr(x), x.shape[0], unit_map, int(active_levels),
                                                 taps, n_taps, _lib.ptr(dout), int(point_stride), int(tap_stride),
                                                 int(level_stride), _lib.ptr(dtable), _lib.current_stream())
    _lib.check(rc, "na_hashgrid_taps_bwd_strided")
    return dtable


class _HashGridTaps(torch.autograd.Function):
    """Autograd wrapper: gradient w.r.t. the table only (numerical-gradient mode never asks for d/dx,
    modules.py:132-175; analytical mode, modules.py:117-130, is outside this path)."""

    @staticmethod
    def forward(ctx, x, table, module, unit_map, active_levels, tap_offsets):
        out = encode_taps(module.desc, table, x, unit_map, active_levels, tap_offsets)
        ctx.save_for_backward(x)
        ctx.module = module
        ctx.args = (unit_map, active_levels, tap_offsets)
        ctx.table_shape = table.shape
        return out

    @staticmethod
    def backward(ctx, dout):
        (x,) = ctx.saved_tensors
        module = ctx.module
        unit_map, active_levels, tap_offsets = ctx.args
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("hash-grid encode: the gradient w.r.t. the query points is not implemented "
                                      "(numerical-gradient mode never needs it) -- detach the points")
        if not ctx.needs_input_grad[1]:
            return None, None, None, None, None, None
        dtable = module._grad_buffer()
        scatter_taps(module.desc, x, unit_map, active_levels, tap_offsets, dout, dtable)
        if module.accumulate_grad_in_place:
            return None, None, None, None, None, None
        return None, dtable, None, None, None, None


class HashGridEncoding(torch.nn.Module):
    """`tcnn.Encoding`-compatible module (attribute `params`, `n_input_dims`, `n_output_dims`).

    accumulate_grad_in_place: when True the backward scatters straight into `params.grad`
    (allocated zero-filled on first use) instead of materialising one 1.46 GB gradient per
    encode call and letting autograd sum them; numerically the same sum, 3x fewer table passes.
    """

    def __init__(self, n_input_dims, encoding_config, seed=1337, dtype=None, accumulate_grad_in_place=False):
        super().__init__()
        if n_input_dims != 3:
            raise NotImplementedError("only 3-D inputs are on the Neuralangelo path")
        cfg = dict(encoding_config)
        otype = cfg.get("otype", "HashGrid")
        if otype not in ("HashGrid", "Grid"):
            raise NotImplementedError(f"encoding otype {otype!r} is not on the Neuralangelo path")
        self.encoding_config = cfg
        self.n_input_dims = n_input_dims
        self.seed = seed
        self.desc = make_desc(cfg.get("n_levels", 16), cfg.get("n_features_per_level", 2),
                              cfg.get("log2_hashmap_size", 19), cfg.get("base_resolution", 16),
                              cfg.get("per_level_scale", 2.0))
        self.n_output_dims = self.desc.n_levels * self.desc.n_features
        self.accumulate_grad_in_place = accumulate_grad_in_place
        self.active_levels = self.desc.n_levels
        n_params = self.desc.total_entries * self.desc.n_features
        gen = torch.Generator(device="cpu").manual_seed(seed)
        # tcnn initialises the grid U(-1e-4, 1e-4); only determinism is needed, not tcnn's RNG stream.
        init = (torch.rand(n_params, generator=gen, dtype=_F32) * 2.0 - 1.0) * 1e-4
        self.params = torch.nn.Parameter(init)
        self.dtype = _F32
        self.identity_map = _lib.UnitMap(0.0, 1.0)

    def _grad_buffer(self):
        if self.accumulate_grad_in_place:
            if self.params.grad is None:
                self.params.grad = torch.zeros_like(self.params)
            return self.params.grad
        return torch.zeros_like(self.params)

    def encode_taps(self, x, tap_offsets, unit_map=None, active_levels=None):
        """[n_taps, N, L*F] features at x + tap_offsets[t] (autograd-aware w.r.t. params)."""
        unit_map = unit_map or self.identity_map
        active_levels = self.active_levels if active_levels is None else active_levels
        tap_offsets = tuple(tuple(float(v) for v in row) for row in tap_offsets)
        x = _as_points(x.reshape(-1, 3))
        if torch.is_grad_enabled() and self.params.requires_grad:
            return _HashGridTaps.apply(x, self.params, self, unit_map, active_levels, tap_offsets)
        return encode_taps(self.desc, self.params.detach(), x, unit_map, active_levels, tap_offsets)

    def forward(self, x):
        return self.encode_taps(x, ((0.0, 0.0, 0.0),))[0]

    def extra_repr(self):
        d = self.desc
        return f"levels={d.n_levels}, features={d.n_features}, entries={d.total_entries}"


def Encoding(n_input_dims, encoding_config, seed=1337, dtype=None):
    """Same factory signature as `tinycudann.Encoding` (call site modules.py:50)."""
    return HashGridEncoding(n_input_dims, encoding_config, seed=seed, dtype=dtype)
