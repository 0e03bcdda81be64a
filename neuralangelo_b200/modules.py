"""Drop-in for `projects/neuralangelo/utils/modules.py`: NeuralSDF, NeuralRGB, BackgroundNeRF with the
reference's constructor arguments, method names, attributes and state-dict keys.

What changes underneath (reference lines in brackets):
  * the hash grid is `hashgrid.HashGridEncoding` (csrc/hashgrid.cu) instead of tinycudann [modules.py:50,85];
    the [0,1] normalisation, the coarse-to-fine mask and the `cat([xyz, enc])` [modules.py:81-94] are folded
    into the encode call: features land at column 3 of a [N,131] row view, inactive levels are written as zeros;
  * compute_gradients in numerical mode evaluates the centre + 4/6 taps with ONE encode launch that shares
    corner rows between taps, followed by one MLP pass over all taps [modules.py:132-175 make 4/6 separate
    encode + MLP round trips].
"""
from functools import partial

import numpy as np
import torch
import torch.nn.functional as F

from . import _lib, hashgrid
from .mlp import MLPforNeuralSDF, MLPwithSkipConnection, get_activation
from .spherical_harmonics import get_spherical_harmonics


def positional_encoding(input, num_freq_bases):
    """nerf_util.py:134-147: [.., N] -> [.., 2*N*L], per input dim: sin(2^l pi x) for all l, then cos."""
    freq = 2 ** torch.arange(num_freq_bases, dtype=torch.float32, device=input.device) * np.pi
    spectrum = input[..., None] * freq
    enc = torch.stack([spectrum.sin(), spectrum.cos()], dim=-2)
    return enc.view(*input.shape[:-1], -1)


_tap_cache = {}


def _tap_tensor(taps, device):
    """Device copy of a tap-offset tuple, cached (a fresh torch.tensor() per call is a pageable H2D copy)."""
    key = (taps, str(device))
    if key not in _tap_cache:
        _tap_cache[key] = torch.tensor(taps, dtype=torch.float32, device=device)
    return _tap_cache[key]


class _EncodeRows(torch.autograd.Function):
    """[T,N,3+L*F] rows = (x + tap_t) ++ hash-grid features of (x + tap_t); gradient to the table only."""

    @staticmethod
    def forward(ctx, pts, table, module, unit_map, active_levels, taps):
        width = module.n_output_dims
        n, n_taps = pts.shape[0], len(taps)
        # rows padded to 3+width+5 floats with the pad in FRONT: features start 32 bytes into a row whose stride is a
        # multiple of 32 bytes (for the default 128 features), so the kernel's whole-row 256-bit stores apply; the
        # returned view drops the pad columns.
        buf = torch.empty(n_taps, n, width + 8, dtype=torch.float32, device=pts.device)
        offs = _tap_tensor(taps, pts.device)
        buf[:, :, 5:8] = pts[None, :, :] + offs[:, None, :]
        hashgrid.encode_taps(module.desc, table, pts, unit_map, active_levels, taps, out=buf[:, :, 8:])
        ctx.save_for_backward(pts)
        ctx.module, ctx.args = module, (unit_map, active_levels, taps)
        return buf[:, :, 5:]

    @staticmethod
    def backward(ctx, drows):
        (pts,) = ctx.saved_tensors
        module = ctx.module
        unit_map, active_levels, taps = ctx.args
        if ctx.needs_input_grad[0]:
            raise NotImplementedError("hash-grid encode: the gradient w.r.t. the query points (analytical normals, "
                                      "pose optimisation; modules.py:117-130) is not implemented -- detach the points")
        if not ctx.needs_input_grad[1]:
            return None, None, None, None, None, None
        dtable = module._grad_buffer()
        hashgrid.scatter_taps(module.desc, pts, unit_map, active_levels, taps, drows[:, :, 3:], dtable)
        if module.accumulate_grad_in_place:
            return None, None, None, None, None, None
        return None, dtable, None, None, None, None


class NeuralSDF(torch.nn.Module):

    # explicit opt-in: evaluate an MLP architecture the fused kernels do not cover with torch ops (never implicit)
    allow_torch_mlp = False

    def __init__(self, cfg_sdf):
        super().__init__()
        self.cfg_sdf = cfg_sdf
        encoding_dim = self.build_encoding(cfg_sdf.encoding)
        self.build_mlp(cfg_sdf.mlp, input_dim=3 + encoding_dim)

    def build_encoding(self, cfg_encoding):
        if cfg_encoding.type == "fourier":
            return 6 * cfg_encoding.levels
        if cfg_encoding.type != "hashgrid":
            raise NotImplementedError("Unknown encoding type")
        # modules.py:36-54
        l_min, l_max = cfg_encoding.hashgrid.min_logres, cfg_encoding.hashgrid.max_logres
        r_min, r_max = 2 ** l_min, 2 ** l_max
        num_levels = cfg_encoding.levels
        self.growth_rate = np.exp((np.log(r_max) - np.log(r_min)) / (num_levels - 1))
        config = dict(otype="HashGrid", n_levels=cfg_encoding.levels, n_features_per_level=cfg_encoding.hashgrid.dim,
                      log2_hashmap_size=cfg_encoding.hashgrid.dict_size,
                      base_resolution=2 ** cfg_encoding.hashgrid.min_logres, per_level_scale=self.growth_rate)
        self.tcnn_encoding = hashgrid.Encoding(3, config)
        self.resolutions = [int(np.floor(r_min * self.growth_rate ** lv).astype(int) + 1) for lv in range(num_levels)]
        vol_min, vol_max = cfg_encoding.hashgrid.range
        self._unit_map = _lib.UnitMap(float(vol_min), float(np.float32(1.0) / np.float32(vol_max - vol_min)))
        self.active_levels = num_levels  # until set_active_levels is called (coarse2fine) -- all levels
        return cfg_encoding.hashgrid.dim * cfg_encoding.levels

    def build_mlp(self, cfg_mlp, input_dim=3):
        layer_dims = [input_dim] + [cfg_mlp.hidden_dim] * cfg_mlp.num_layers + [cfg_mlp.hidden_dim]
        activ = get_activation(cfg_mlp.activ, **cfg_mlp.activ_params)
        self.mlp = MLPforNeuralSDF(layer_dims, skip_connection=cfg_mlp.skip, activ=activ,
                                   use_weightnorm=cfg_mlp.weight_norm, geometric_init=cfg_mlp.geometric_init,
                                   out_bias=cfg_mlp.out_bias, invert=cfg_mlp.inside_out)

    # ------------------------------------------------------------------------------------------
    def _hashgrid(self):
        return self.cfg_sdf.encoding.type == "hashgrid"

    def _levels_in_use(self):
        c2f = self.cfg_sdf.encoding.coarse2fine
        return self.active_levels if c2f.enabled else self.cfg_sdf.encoding.levels

    def encode_rows(self, points_flat, taps=((0.0, 0.0, 0.0),)):
        """[T,N,3+LD]: xyz ++ masked features at points + taps (one launch)."""
        enc = self.tcnn_encoding
        pts = points_flat.float().contiguous()
        if torch.is_grad_enabled() and enc.params.requires_grad:
            return _EncodeRows.apply(pts, enc.params, enc, self._unit_map, self._levels_in_use(), tuple(taps))
        return _EncodeRows.forward(_NoCtx(), pts, enc.params.detach(), enc, self._unit_map, self._levels_in_use(),
                                   tuple(taps))

    # ---- chunk planes: the layout both the hash-grid kernels and the SDF-only MLP kernel are fastest in ------------
    def _planes_supported(self):
        """8 features per level (one 32-byte chunk per level and point) and a level count the pipelined MLP kernel
        takes; otherwise the row layout is used."""
        if not self._hashgrid():
            return False
        desc = self.tcnn_encoding.desc
        if desc.n_features != 8 or desc.n_levels % 4 != 0 or desc.n_levels > 16:
            return False
        packed = self._packed_mlp() if self.tcnn_encoding.params.is_cuda else None
        return packed is not None and packed.n in (128, 256)

    def encode_planes(self, points_flat, taps=((0.0, 0.0, 0.0),), table=None):
        """[L+1, T*N, 8] chunk planes: plane l < L holds level l's features of point n + tap t at row t*N + n
        (zeros for levels masked by coarse-to-fine), plane L holds xyz of that row in its first three floats."""
        enc = self.tcnn_encoding
        pts = points_flat.float().contiguous()
        n, n_taps, n_levels = pts.shape[0], len(taps), enc.desc.n_levels
        planes = torch.empty(n_levels + 1, n_taps * n, 8, dtype=torch.float32, device=pts.device)
        hashgrid.encode_taps_strided(enc.desc, enc.params.detach() if table is None else table, pts, self._unit_map,
                                     self._levels_in_use(), taps, planes, 8, n * 8, n_taps * n * 8)
        xyz = planes[n_levels].view(n_taps, n, 8)
        xyz[:, :, :3] = pts[None, :, :] + _tap_tensor(tuple(taps), pts.device)[:, None, :]
        return planes

    def sdf_planes(self, points_flat, taps=((0.0, 0.0, 0.0),)):
        """SDF at points + taps without an autograd graph: one encode launch into chunk planes, one launch of the
        pipelined SDF-only MLP.  Returns [T, N, 1]."""
        planes = self.encode_planes(points_flat, taps)
        sdf = self._packed_mlp().forward_planes(planes, planes.shape[1])
        return sdf.view(len(taps), points_flat.shape[0], 1)

    def _lattice_fused_supported(self):
        enc, mlp = self.tcnn_encoding, self._packed_mlp()
        return (mlp is not None and enc.desc.n_levels == 16 and enc.desc.n_features == 8 and mlp.n == 256
                and mlp.k0 == 131 and mlp.k0_pad == 144)

    def sdf_lattice(self, gx, gy, gz, fused=None, scale=1.0):
        """SDF on the lattice gx x gy x gz (float32 CUDA axis arrays, x slowest: one block of extract_mesh) without an
        [N,3] coordinate tensor.  Default shape (16 levels x 8 features, 256-wide MLP): ONE kernel (na_lattice_sdf) that
        gathers along z runs straight into the tensor-core operand; other shapes (or fused=False): the encode kernel
        derives each point from the axes and writes the MLP's chunk planes (na_hashgrid_lattice_fwd_strided), then one
        launch of the pipelined SDF-only MLP.  Returns scale * sdf as [nx,ny,nz] (scale = -1: the flip of
        scripts/extract_mesh.py:85, applied in the kernel's last store)."""
        import ctypes
        enc = self.tcnn_encoding
        if not self._planes_supported():
            raise NotImplementedError("sdf_lattice needs the chunk-plane path (8 features per level, fused MLP envelope)")
        nx, ny, nz = gx.numel(), gy.numel(), gz.numel()
        n, n_levels = nx * ny * nz, enc.desc.n_levels
        if fused is None:
            fused = self._lattice_fused_supported()
        if fused:
            mlp = self._packed_mlp()
            w_hi, w_lo, head = mlp.lattice_operands()
            sdf = torch.empty(nx, ny, nz, dtype=torch.float32, device=gx.device)
            rc = _lib.lib().na_lattice_sdf(
                ctypes.byref(enc.desc), _lib.ptr(enc.params.detach()), _lib.ptr(gx), _lib.ptr(gy), _lib.ptr(gz), nx, ny, nz,
                self._unit_map, int(self._levels_in_use()), _lib.ptr(w_hi), _lib.ptr(w_lo), mlp.k0_pad, mlp.n,
                _lib.ptr(head), _lib.ptr(mlp.b_head), float(scale), _lib.ptr(sdf), _lib.current_stream())
            _lib.check(rc, "na_lattice_sdf")
            return sdf
        planes = torch.empty(n_levels + 1, n, 8, dtype=torch.float32, device=gx.device)
        rc = _lib.lib().na_hashgrid_lattice_fwd_strided(
            ctypes.byref(enc.desc), _lib.ptr(enc.params.detach()), _lib.ptr(gx), _lib.ptr(gy), _lib.ptr(gz), nx, ny, nz,
            self._unit_map, int(self._levels_in_use()), _lib.ptr(planes), 8, n * 8, _lib.ptr(planes[n_levels]), 8,
            _lib.current_stream())
        _lib.check(rc, "na_hashgrid_lattice_fwd_strided")
        sdf = self._packed_mlp().forward_planes(planes, n).view(nx, ny, nz)
        return sdf if scale == 1.0 else sdf * scale

    def _needs_graph(self):
        return torch.is_grad_enabled() and (self.tcnn_encoding.params.requires_grad
                                            or any(p.requires_grad for p in self.mlp.parameters()))

    def forward(self, points_3D, with_sdf=True, with_feat=True):
        if with_sdf and not with_feat and points_3D.is_cuda and not (torch.is_grad_enabled() and points_3D.requires_grad) \
                and self._planes_supported() and not self._needs_graph():
            return self.sdf_planes(points_3D.reshape(-1, 3))[0].view(*points_3D.shape[:-1], 1), None
        points_enc = self.encode(points_3D)  # [...,3+LD]
        sdf, feat = self.run_mlp(points_enc, with_sdf=with_sdf, with_feat=with_feat)
        return sdf, feat  # [...,1],[...,K]

    # ---- MLP dispatch: tcgen05 fused kernel when no autograd graph is needed, torch ops otherwise ----------
    def invalidate_packed_weights(self):
        """Call after the MLP weights changed through a path torch cannot see (the fused optimizer kernel)."""
        self._weights_epoch = getattr(self, "_weights_epoch", 0) + 1

    def _packed_mlp(self):
        from .fused_mlp import PackedSdfMlp
        key = (getattr(self, "_weights_epoch", 0),) + tuple(p._version for p in self.mlp.parameters()) + \
            (next(self.mlp.parameters()).device,)
        if getattr(self, "_packed_key", None) != key:
            old = getattr(self, "_packed", None)
            if old is not None and old.b0.device == key[-1]:
                old.update_(self.mlp)       # in place: addresses recorded in a CUDA graph stay valid
            else:
                try:
                    self._packed = PackedSdfMlp(self.mlp, beta=self.cfg_sdf.mlp.activ_params.get("beta", 1.0)) \
                        if self.cfg_sdf.mlp.activ == "softplus" else None
                except NotImplementedError:
                    self._packed = None  # architecture outside the fused kernel's envelope: torch ops
            self._packed_key = key
        return self._packed

    def refresh_packed_weights(self):
        """Re-pack now (eagerly).  The training step calls this right after the optimizer so that the packing
        never happens inside a captured graph, where it would be frozen."""
        self.invalidate_packed_weights()
        if next(self.mlp.parameters()).is_cuda:
            self._packed_mlp()

    def run_mlp(self, rows, with_sdf=True, with_feat=True):
        """rows [...,3+LD] -> [sdf [...,1] | None, feat [...,K] | None]."""
        needs_graph = torch.is_grad_enabled() and (rows.requires_grad or any(p.requires_grad for p in self.mlp.parameters()))
        _lib.require_cuda(rows)
        packed = self._packed_mlp()
        if packed is None:
            # outside the tcgen05 kernels' envelope (softplus, one hidden layer, no skip, width 64..256, input <= 256):
            # there is no silent fallback -- torch ops run only when the caller has opted in explicitly
            if not self.allow_torch_mlp:
                raise NotImplementedError(
                    "NeuralSDF MLP is outside the fused kernel's envelope (softplus, num_layers=1, no skip connection, "
                    "hidden width 64/128/192/256, input width <= 256); set NeuralSDF.allow_torch_mlp = True to "
                    "evaluate it with torch ops")
            return self.mlp(rows, with_sdf=with_sdf, with_feat=with_feat)
        flat = rows.reshape(-1, rows.shape[-1])
        if flat.stride(-1) != 1 or flat.dtype != torch.float32:
            flat = flat.float().contiguous()
        if needs_graph:   # tcgen05 forward + input-gradient chain, weight gradients through torch (fused_mlp.py)
            from .fused_mlp import fused_sdf_mlp
            sdf, feat = fused_sdf_mlp(self.mlp, flat, packed.beta, with_sdf=with_sdf, with_feat=with_feat, packed=packed)
        else:
            sdf, feat = packed.forward(flat, with_sdf=with_sdf, with_feat=with_feat)
        lead = rows.shape[:-1]
        return [sdf.view(*lead, 1) if sdf is not None else None, feat.view(*lead, -1) if feat is not None else None]

    def sdf(self, points_3D):
        return self.forward(points_3D, with_sdf=True, with_feat=False)[0]

    def encode(self, points_3D):
        if self._hashgrid():
            rows = self.encode_rows(points_3D.reshape(-1, 3))[0]
            return rows.view(*points_3D.shape[:-1], rows.shape[-1])
        if self.cfg_sdf.encoding.type != "fourier":
            raise NotImplementedError("Unknown encoding type")
        points_enc = positional_encoding(points_3D, num_freq_bases=self.cfg_sdf.encoding.levels)
        if self.cfg_sdf.encoding.coarse2fine.enabled:
            mask = torch.zeros_like(points_enc)
            mask[..., :(self.active_levels * 6)] = 1
            points_enc = points_enc * mask
        return torch.cat([points_3D, points_enc], dim=-1)

    def set_active_levels(self, current_iter=None):
        anneal_levels = max((current_iter - self.warm_up_end) // self.cfg_sdf.encoding.coarse2fine.step, 1)
        self.anneal_levels = min(self.cfg_sdf.encoding.levels, anneal_levels)
        self.active_levels = max(self.cfg_sdf.encoding.coarse2fine.init_active_level, self.anneal_levels)

    def set_normal_epsilon(self):
        if self.cfg_sdf.encoding.coarse2fine.enabled:
            epsilon_res = self.resolutions[self.anneal_levels - 1]
        else:
            epsilon_res = self.resolutions[-1]
        self.normal_eps = 1. / epsilon_res

    def _tap_offsets(self):
        taps = self.cfg_sdf.gradient.taps
        if taps == 6:
            e = float(np.float32(self.normal_eps))  # torch.tensor([eps,0,0], dtype=float32), modules.py:135-137
            return e, ((e, 0, 0), (-e, 0, 0), (0, e, 0), (0, -e, 0), (0, 0, e), (0, 0, -e))
        if taps == 4:
            eps = self.normal_eps / np.sqrt(3)  # modules.py:158
            e = float(np.float32(eps))          # k * eps is evaluated in float32 (k is a float32 tensor)
            return eps, ((e, -e, -e), (-e, -e, e), (-e, e, -e), (e, e, e))
        raise ValueError("Only support 4 or 6 taps.")

    def compute_gradients(self, x, training=False, sdf=None):
        """modules.py:115-178.  Note: hessian is the diagonal (6 taps) or trace/3 replicated (4 taps)."""
        mode = self.cfg_sdf.gradient.mode
        if mode == "analytical":
            raise NotImplementedError("analytical gradients need the encode's second derivative; the shipped "
                                      "default (base.yaml:67-68) is numerical")
        if mode != "numerical":
            raise ValueError(mode)
        eps, offsets = self._tap_offsets()
        taps_sdf = None
        if self._hashgrid() and x.is_cuda and self._planes_supported() and not self._needs_graph() \
                and not (torch.is_grad_enabled() and x.requires_grad):
            taps_sdf = self.sdf_planes(x.reshape(-1, 3), taps=offsets)                # [T,N,1]
        elif self._hashgrid():
            rows = self.encode_rows(x.reshape(-1, 3), taps=offsets)                 # [T,N,3+LD], one encode launch
            taps_sdf = self.run_mlp(rows, with_sdf=True, with_feat=False)[0]         # [T,N,1], one MLP pass
        if taps_sdf is not None and x.is_cuda:
            # gradient / curvature arithmetic as ONE launch per direction (csrc/glue.cu) instead of ~15 + ~30 ATen kernels
            from . import glue
            return glue.tap_combine(taps_sdf.view(len(offsets), *x.shape[:-1], 1), sdf, self.cfg_sdf.gradient.taps, eps,
                                    training)
        if taps_sdf is not None:
            s = [t.view(*x.shape[:-1], 1) for t in taps_sdf.unbind(0)]
        else:
            s = [self.sdf(x + torch.tensor(o, dtype=x.dtype, device=x.device)) for o in offsets]
        hessian = None
        if self.cfg_sdf.gradient.taps == 6:
            gradient = torch.cat([(s[0] - s[1]) / (2 * eps), (s[2] - s[3]) / (2 * eps), (s[4] - s[5]) / (2 * eps)], dim=-1)
            if training:
                assert sdf is not None
                hessian = torch.cat([(s[0] + s[1] - 2 * sdf) / (eps ** 2), (s[2] + s[3] - 2 * sdf) / (eps ** 2),
                                     (s[4] + s[5] - 2 * sdf) / (eps ** 2)], dim=-1)
        else:
            k = _tap_tensor(((1., -1., -1.), (-1., -1., 1.), (-1., 1., -1.), (1., 1., 1.)), x.device).to(x.dtype)
            gradient = (k[0] * s[0] + k[1] * s[1] + k[2] * s[2] + k[3] * s[3]) / (4.0 * eps)
            if training:
                assert sdf is not None
                hessian_xx = ((s[0] + s[1] + s[2] + s[3]) / 2.0 - 2 * sdf) / eps ** 2
                hessian = torch.cat([hessian_xx, hessian_xx, hessian_xx], dim=-1) / 3.0
        return gradient, hessian


# The colour / background networks run on the tcgen05 chain kernel (csrc/mlp_chain.cu) whenever no autograd graph is
# needed (validation renders, texture extraction): one launch per network.  With a graph (training) the same kernel can
# run the forward (saving the layer activations) and, as a second stage program with relu-mask epilogues, the
# input-gradient chain (fused_chain.FusedMlpFunction) -- parity-tested, but measured SLOWER than the library path on
# B200 (default config, 512 rays: forward+backward 6.85 ms against 5.14 ms with TF32 library GEMMs, plus 1.8 ms per
# step of operand re-packing; profiles/r02_training.md): the chain kernel runs its stages one after the other on a
# single accumulator and pays three fp16 MMAs per product.  So it stays opt-in; the weight and bias gradients of these
# networks run on the tcgen05 weight-gradient kernel either way (mlp.linear_wgrad).
FUSED_CHAIN_IN_TRAINING = False


def _chain_supported(mlp):
    from . import fused_chain
    return fused_chain._supported(mlp)


class _NoCtx:
    """Stand-in for the autograd context when the encode runs without gradient tracking."""
    needs_input_grad = (False,) * 6

    def save_for_backward(self, *a):
        pass


class NeuralRGB(torch.nn.Module):

    def __init__(self, cfg_rgb, feat_dim, appear_embed):
        super().__init__()
        self.cfg_rgb = cfg_rgb
        self.cfg_appear_embed = appear_embed
        encoding_view_dim = self.build_encoding(cfg_rgb.encoding_view)
        input_base_dim = 6 if cfg_rgb.mode == "idr" else 3
        input_dim = input_base_dim + encoding_view_dim + feat_dim + (appear_embed.dim if appear_embed.enabled else 0)
        self.build_mlp(cfg_rgb.mlp, input_dim=input_dim)

    def build_encoding(self, cfg_encoding_view):
        if cfg_encoding_view.type == "fourier":
            return 6 * cfg_encoding_view.levels
        if cfg_encoding_view.type == "spherical":
            self.spherical_harmonic_encoding = partial(get_spherical_harmonics, levels=cfg_encoding_view.levels)
            return (cfg_encoding_view.levels + 1) ** 2
        raise NotImplementedError("Unknown encoding type")

    def build_mlp(self, cfg_mlp, input_dim=3):
        layer_dims = [input_dim] + [cfg_mlp.hidden_dim] * cfg_mlp.num_layers + [3]
        activ = get_activation(cfg_mlp.activ, **cfg_mlp.activ_params)
        self.mlp = MLPwithSkipConnection(layer_dims, skip_connection=cfg_mlp.skip, activ=activ,
                                         use_weightnorm=cfg_mlp.weight_norm)

    def forward(self, points_3D, normals, rays_unit, feats, app):
        view_enc = self.encode_view(rays_unit)
        input_list = [points_3D, view_enc, normals, feats]
        if app is not None:
            input_list.append(app)
        if self.cfg_rgb.mode == "no_view_dir":
            input_list = [t for t in input_list if t is not view_enc]
        if self.cfg_rgb.mode == "no_normal":
            input_list = [t for t in input_list if t is not normals]
        input_vec = torch.cat(input_list, dim=-1)
        return self._run_mlp(input_vec)

    def _run_mlp(self, input_vec):
        if input_vec.is_cuda and _chain_supported(self.mlp):
            # tcgen05 chain kernel (csrc/mlp_chain.cu): all five layers + sigmoid in one launch
            from . import fused_chain
            k_in = self.mlp.linears[0].weight_shape()[1] if hasattr(self.mlp.linears[0], "weight_shape") else \
                self.mlp.linears[0].in_features
            flat = input_vec.reshape(-1, input_vec.shape[-1])[:, :k_in].float().contiguous()   # without any zero padding
            needs_graph = torch.is_grad_enabled() and (flat.requires_grad or any(p.requires_grad for p in self.mlp.parameters()))
            if not needs_graph:
                out = fused_chain.mlp_forward(self.mlp, flat, final_act="sigmoid")
                return out.view(*input_vec.shape[:-1], out.shape[-1])
            if FUSED_CHAIN_IN_TRAINING:
                out = fused_chain.fused_mlp(self.mlp, flat, final_act="sigmoid")
                return out.view(*input_vec.shape[:-1], out.shape[-1])
        return self.mlp(input_vec).sigmoid()

    def encode_view(self, rays_unit):
        if self.cfg_rgb.encoding_view.type == "fourier":
            return positional_encoding(rays_unit, num_freq_bases=self.cfg_rgb.encoding_view.levels)
        if self.cfg_rgb.encoding_view.type == "spherical":
            return self.spherical_harmonic_encoding(rays_unit)
        raise NotImplementedError("Unknown encoding type")

    def forward_from_gradients(self, points_3D, gradients, ray_unit, feats, app):
        """forward(points, normalize(gradients), ray_unit broadcast over the samples, feats, app) as the reference's
        model.py:179-183 calls it, with the input row built by ONE kernel (csrc/glue.cu: SH of the view direction,
        normalisation and the 278-wide concatenation) where the configuration allows (idr mode, spherical view encoding,
        CUDA); ray_unit is [B,R,3]."""
        if self.cfg_rgb.mode == "idr" and self.cfg_rgb.encoding_view.type == "spherical" and feats.is_cuda and \
                self.cfg_rgb.encoding_view.levels <= 4:
            from . import glue
            # with a graph to build the rows go to library GEMMs: K padded to a multiple of four (see mlp.linear)
            pad = 4 if (torch.is_grad_enabled() and not FUSED_CHAIN_IN_TRAINING and not self.mlp.skip_connection) else 1
            input_vec = glue.rgb_input(points_3D, ray_unit, gradients, feats, app, self.cfg_rgb.encoding_view.levels,
                                       pad_to=pad)
            return self._run_mlp(input_vec)
        normals = F.normalize(gradients, dim=-1)
        rays_unit = ray_unit[..., None, :].expand_as(points_3D).contiguous()
        return self.forward(points_3D, normals, rays_unit, feats, app)


class BackgroundNeRF(torch.nn.Module):

    def __init__(self, cfg_background, appear_embed):
        super().__init__()
        self.cfg_background = cfg_background
        self.cfg_appear_embed = appear_embed
        encoding_dim, encoding_view_dim = self.build_encoding(cfg_background.encoding, cfg_background.encoding_view)
        input_dim = 4 + encoding_dim
        input_view_dim = cfg_background.mlp.hidden_dim + encoding_view_dim + \
            (appear_embed.dim if appear_embed.enabled else 0)
        self.build_mlp(cfg_background.mlp, input_dim=input_dim, input_view_dim=input_view_dim)

    def build_encoding(self, cfg_encoding, cfg_encoding_view):
        if cfg_encoding.type != "fourier":
            raise NotImplementedError("Unknown encoding type")
        encoding_dim = 8 * cfg_encoding.levels
        if cfg_encoding_view.type == "fourier":
            encoding_view_dim = 6 * cfg_encoding_view.levels
        elif cfg_encoding_view.type == "spherical":
            self.spherical_harmonic_encoding = partial(get_spherical_harmonics, levels=cfg_encoding_view.levels)
            encoding_view_dim = (cfg_encoding_view.levels + 1) ** 2
        else:
            raise NotImplementedError("Unknown encoding type")
        return encoding_dim, encoding_view_dim

    def build_mlp(self, cfg_mlp, input_dim=3, input_view_dim=3):
        activ = get_activation(cfg_mlp.activ, **cfg_mlp.activ_params)
        layer_dims = [input_dim] + [cfg_mlp.hidden_dim] * (cfg_mlp.num_layers - 1) + [cfg_mlp.hidden_dim + 1]
        self.mlp_feat = MLPwithSkipConnection(layer_dims, skip_connection=cfg_mlp.skip, activ=activ)
        self.activ_density = get_activation(cfg_mlp.activ_density, **cfg_mlp.activ_density_params)
        layer_dims_rgb = [input_view_dim] + [cfg_mlp.hidden_dim_rgb] * (cfg_mlp.num_layers_rgb - 1) + [3]
        self.mlp_rgb = MLPwithSkipConnection(layer_dims_rgb, skip_connection=cfg_mlp.skip_rgb, activ=activ)

    def forward(self, points_3D, rays_unit, app_outside):
        points_enc = self.encode(points_3D)
        fused = points_enc.is_cuda and app_outside is None and self.cfg_background.view_dep and \
            _chain_supported(self.mlp_feat) and _chain_supported(self.mlp_rgb)
        if fused:
            from . import fused_chain
            lead = points_enc.shape[:-1]
            enc = points_enc.reshape(-1, points_enc.shape[-1]).float().contiguous()
            venc = self.encode_view(rays_unit).reshape(enc.shape[0], -1).float().contiguous()
            needs_graph = torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters())
            if not needs_graph:   # whole network (both MLPs, density, sigmoid) in ONE launch
                rgb, density = fused_chain.background_forward(self, enc, venc)
                return rgb.view(*lead, 3), density.view(*lead)
            if FUSED_CHAIN_IN_TRAINING:
                out = fused_chain.fused_mlp(self.mlp_feat, enc, final_act="none")
                density, feat = self.activ_density(out[..., 0]), self.mlp_feat.activ(out[..., 1:])
                rgb = fused_chain.fused_mlp(self.mlp_rgb, torch.cat([feat, venc], dim=-1), final_act="sigmoid")
                return rgb.view(*lead, 3), density.view(*lead)
        out = self.mlp_feat(points_enc)
        density, feat = self.activ_density(out[..., 0]), self.mlp_feat.activ(out[..., 1:])
        if not self.cfg_background.view_dep:
            raise NotImplementedError
        input_list = [feat, self.encode_view(rays_unit)]
        if app_outside is not None:
            input_list.append(app_outside)
        rgb = self.mlp_rgb(torch.cat(input_list, dim=-1)).sigmoid()
        return rgb, density

    def encode(self, points_3D):
        # inverted-sphere reparametrisation (x/|x|, 1/|x|), modules.py:289-292
        norm = points_3D.norm(dim=-1, keepdim=True)
        points = torch.cat([points_3D / norm, 1.0 / norm], dim=-1)
        if self.cfg_background.encoding.type != "fourier":
            raise NotImplementedError("Unknown encoding type")
        return torch.cat([points, positional_encoding(points, num_freq_bases=self.cfg_background.encoding.levels)], dim=-1)

    def encode_view(self, rays_unit):
        if self.cfg_background.encoding_view.type == "fourier":
            return positional_encoding(rays_unit, num_freq_bases=self.cfg_background.encoding_view.levels)
        if self.cfg_background.encoding_view.type == "spherical":
            return self.spherical_harmonic_encoding(rays_unit)
        raise NotImplementedError("Unknown encoding type")
