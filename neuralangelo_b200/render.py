"""Drop-in for `projects/nerf/utils/render.py` (the three functions on the Neuralangelo path) plus the
NeuS-specific pieces of `projects/neuralangelo/model.py`, all running in libna_b200.so (csrc/render.cu).

Same names, argument meaning and tensor shapes as the reference:
    volume_rendering_alphas_dist(densities [B,R,N], dists [B,R,N,1], dist_far [B,R,1,1]|None) -> [B,R,N]
    alpha_compositing_weights(alphas [B,R,N]) -> [B,R,N,1]
    composite(quantities [B,R,N,K] | float, weights [B,R,N,1]) -> [B,R,K]
and, beyond the reference's module, the fused forms its Model methods reduce to:
    neus_alphas(...)   == Model.compute_neus_alphas (model.py:275-292)
    neus_render(...)   == alphas(object) ++ alphas(background) -> weights -> composite (model.py:132-141)
Every function is differentiable through hand-written backward kernels.
"""
import torch

from . import _lib
from ._lib import f32c, ptr

_L = _lib.lib


def _rays(t, trailing):
    """[B,R,N,*trailing] -> ([B*R,N,*trailing] contiguous fp32, (B,R,N))."""
    shape = t.shape
    lead = shape[:len(shape) - 1 - len(trailing)] if trailing else shape[:-1]
    n = shape[len(lead)]
    flat = f32c(t).reshape(-1, n, *trailing)
    return flat, lead, n


class _AlphaWeights(torch.autograd.Function):
    @staticmethod
    def forward(ctx, alphas):
        _lib.require_cuda(alphas)
        a = f32c(alphas)
        n = a.shape[-1]
        rays = a.numel() // n
        w = torch.empty_like(a)
        _lib.check(_L().na_alpha_weights_fwd(ptr(a), rays, n, ptr(w), _lib.current_stream()), "na_alpha_weights_fwd")
        ctx.save_for_backward(a)
        return w

    @staticmethod
    def backward(ctx, dw):
        (a,) = ctx.saved_tensors
        n = a.shape[-1]
        rays = a.numel() // n
        dw = f32c(dw)
        da = torch.empty_like(a)
        _lib.check(_L().na_alpha_weights_bwd(ptr(a), ptr(dw), rays, n, ptr(da), _lib.current_stream()),
                   "na_alpha_weights_bwd")
        return da


def alpha_compositing_weights(alphas):
    """render.py:87-99.  alphas [B,R,N] -> weights [B,R,N,1]."""
    return _AlphaWeights.apply(alphas)[..., None]


class _Composite(torch.autograd.Function):
    @staticmethod
    def forward(ctx, q, w):
        _lib.require_cuda(q, w)
        q, w = f32c(q), f32c(w)
        k, n = q.shape[-1], q.shape[-2]
        rays = w.numel() // n
        out = torch.empty(*q.shape[:-2], k, dtype=torch.float32, device=q.device)
        _lib.check(_L().na_composite_fwd(ptr(q), ptr(w), rays, n, k, ptr(out), _lib.current_stream()), "na_composite_fwd")
        ctx.save_for_backward(q, w)
        return out

    @staticmethod
    def backward(ctx, d_out):
        q, w = ctx.saved_tensors
        k, n = q.shape[-1], q.shape[-2]
        rays = w.numel() // n
        d_out = f32c(d_out)
        dq = torch.empty_like(q) if ctx.needs_input_grad[0] else None
        dw = torch.empty_like(w) if ctx.needs_input_grad[1] else None
        _lib.check(_L().na_composite_bwd(ptr(q), ptr(w), ptr(d_out), rays, n, k, ptr(dq), ptr(dw), _lib.current_stream()),
                   "na_composite_bwd")
        return dq, dw


def composite(quantities, weights):
    """render.py:102-112.  quantities [B,R,N,K] (or a python scalar, as model.py:138,171 pass 1.),
    weights [B,R,N,1] -> [B,R,K]."""
    w = weights[..., 0]
    if not torch.is_tensor(quantities):
        quantities = torch.full((*w.shape, 1), float(quantities), dtype=torch.float32, device=w.device)
    return _Composite.apply(quantities, w)


class _DensityAlphas(torch.autograd.Function):
    @staticmethod
    def forward(ctx, dens, dists, dist_far):
        _lib.require_cuda(dens, dists)
        dens, dists, dist_far = f32c(dens), f32c(dists), f32c(dist_far)
        n = dens.shape[-1]
        rays = dens.numel() // n
        out = torch.empty_like(dens)
        _lib.check(_L().na_density_alphas_fwd(ptr(dens), ptr(dists), ptr(dist_far), rays, n, ptr(out),
                                              _lib.current_stream()), "na_density_alphas_fwd")
        ctx.save_for_backward(dens, dists, dist_far)
        return out

    @staticmethod
    def backward(ctx, d_alphas):
        dens, dists, dist_far = ctx.saved_tensors
        n = dens.shape[-1]
        rays = dens.numel() // n
        d_dens = torch.empty_like(dens)
        _lib.check(_L().na_density_alphas_bwd(ptr(dens), ptr(dists), ptr(dist_far), ptr(f32c(d_alphas)), rays, n,
                                              ptr(d_dens), _lib.current_stream()), "na_density_alphas_bwd")
        return d_dens, None, None


def volume_rendering_alphas_dist(densities, dists, dist_far=None):
    """render.py:68-84.  densities [B,R,N], dists [B,R,N,1], dist_far [B,R,1,1]|None -> alphas [B,R,N].
    Distances are sampling positions (produced under no_grad at model.py:193-194): no gradient flows to them."""
    far = None if dist_far is None else dist_far.reshape(densities.shape[:-1])
    return _DensityAlphas.apply(densities, dists[..., 0].detach(), far)


def _anneal_tensor(progress, anneal_end, anneal_ratio, device):
    """Device scalar for min(progress / anneal_end, 1): either the caller's persistent tensor (kept up to date
    outside a captured graph) or a fresh one filled from the python value."""
    if anneal_ratio is not None:
        return anneal_ratio
    return torch.full((1,), min(progress / anneal_end, 1.), dtype=torch.float32, device=device)


class _NeusAlphas(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sdf, grad, s_var, dists, ray_unit, dist_far, anneal_ratio):
        _lib.require_cuda(sdf, grad, s_var)
        sdf, grad, dists, ray_unit, dist_far = f32c(sdf), f32c(grad), f32c(dists), f32c(ray_unit), f32c(dist_far)
        s_var = f32c(s_var)
        n = sdf.shape[-1]
        rays = sdf.numel() // n
        out = torch.empty_like(sdf)
        _lib.check(_L().na_neus_alphas_fwd(ptr(sdf), ptr(grad), ptr(dists), ptr(ray_unit), ptr(dist_far), ptr(s_var),
                                           ptr(anneal_ratio), rays, n, ptr(out), _lib.current_stream()),
                   "na_neus_alphas_fwd")
        ctx.save_for_backward(sdf, grad, dists, ray_unit, dist_far, s_var)
        ctx.anneal_ratio = anneal_ratio
        return out

    @staticmethod
    def backward(ctx, d_alphas):
        sdf, grad, dists, ray_unit, dist_far, s_var = ctx.saved_tensors
        n = sdf.shape[-1]
        rays = sdf.numel() // n
        d_sdf, d_grad = torch.empty_like(sdf), torch.empty_like(grad)
        d_s = torch.zeros_like(s_var)
        _lib.check(_L().na_neus_alphas_bwd(ptr(sdf), ptr(grad), ptr(dists), ptr(ray_unit), ptr(dist_far), ptr(s_var),
                                           ptr(ctx.anneal_ratio), ptr(f32c(d_alphas)), rays, n, ptr(d_sdf), ptr(d_grad),
                                           ptr(d_s), _lib.current_stream()), "na_neus_alphas_bwd")
        return d_sdf, d_grad, d_s, None, None, None, None


def neus_alphas(ray_unit, sdfs, gradients, dists, s_var, dist_far=None, progress=1., anneal_end=0.1, anneal_ratio=None):
    """Model.compute_neus_alphas (model.py:275-292) with _get_iter_cos (:294-298).
    ray_unit [B,R,3], sdfs [B,R,N,1], gradients [B,R,N,3], dists [B,R,N,1], dist_far [B,R,1,1]|None -> [B,R,N]."""
    anneal_ratio = _anneal_tensor(progress, anneal_end, anneal_ratio, sdfs.device)
    far = None if dist_far is None else dist_far.reshape(sdfs.shape[:-2])
    return _NeusAlphas.apply(sdfs[..., 0], gradients, s_var, dists[..., 0].detach(), ray_unit.detach(), far,
                             anneal_ratio)


class _NeusRender(torch.autograd.Function):
    @staticmethod
    def forward(ctx, sdf, grad, s_var, rgb_obj, dens_bg, rgb_bg, dists_obj, ray_unit, dist_far, dists_bg,
                anneal_ratio, want_weights):
        _lib.require_cuda(sdf, grad, s_var, rgb_obj)
        sdf, grad, rgb_obj, dists_obj, ray_unit, dist_far = map(f32c, (sdf, grad, rgb_obj, dists_obj, ray_unit, dist_far))
        dens_bg, rgb_bg, dists_bg, s_var = map(f32c, (dens_bg, rgb_bg, dists_bg, s_var))
        n_obj = sdf.shape[-1]
        n_bg = 0 if dens_bg is None else dens_bg.shape[-1]
        lead = sdf.shape[:-1]
        rays = sdf.numel() // n_obj
        rgb = torch.empty(*lead, 3, dtype=torch.float32, device=sdf.device)
        weights = torch.empty(*lead, n_obj + n_bg, dtype=torch.float32, device=sdf.device) if want_weights else None
        opacity = torch.empty(*lead, dtype=torch.float32, device=sdf.device)
        _lib.check(_L().na_neus_render_fwd(ptr(sdf), ptr(grad), ptr(dists_obj), ptr(ray_unit), ptr(dist_far), ptr(s_var),
                                           ptr(anneal_ratio), ptr(rgb_obj), n_obj, ptr(dens_bg), ptr(dists_bg),
                                           ptr(rgb_bg), n_bg, rays, ptr(rgb), ptr(weights), ptr(opacity),
                                           _lib.current_stream()), "na_neus_render_fwd")
        ctx.save_for_backward(sdf, grad, s_var, rgb_obj, dens_bg, rgb_bg, dists_obj, ray_unit, dist_far, dists_bg)
        ctx.anneal_ratio = anneal_ratio
        # opacity = sum of the weights is differentiable (model.py:138-140 adds 1 - opacity_all to the colour for a
        # white background); unused outputs arrive as None in backward instead of materialised zeros
        ctx.set_materialize_grads(False)
        if weights is None:
            weights = torch.empty(0, device=sdf.device)
            ctx.mark_non_differentiable(weights)
        return rgb, weights, opacity

    @staticmethod
    def backward(ctx, d_rgb, d_weights, d_opacity):
        sdf, grad, s_var, rgb_obj, dens_bg, rgb_bg, dists_obj, ray_unit, dist_far, dists_bg = ctx.saved_tensors
        n_obj = sdf.shape[-1]
        n_bg = 0 if dens_bg is None else dens_bg.shape[-1]
        rays = sdf.numel() // n_obj
        if d_rgb is None:
            d_rgb = torch.zeros(*sdf.shape[:-1], 3, dtype=torch.float32, device=sdf.device)
        d_sdf, d_grad, d_rgb_obj = torch.empty_like(sdf), torch.empty_like(grad), torch.empty_like(rgb_obj)
        d_s = torch.zeros_like(s_var)
        d_dens = torch.empty_like(dens_bg) if n_bg else None
        d_rgb_bg = torch.empty_like(rgb_bg) if n_bg else None
        dw = f32c(d_weights) if (d_weights is not None and d_weights.numel() == rays * (n_obj + n_bg)) else None
        if d_opacity is not None:   # opacity = sum_i w_i  =>  every weight receives d_opacity of its ray
            d_op = f32c(d_opacity).reshape(*sdf.shape[:-1], 1)
            dw = (dw.reshape(*sdf.shape[:-1], n_obj + n_bg) + d_op) if dw is not None else \
                d_op.expand(*sdf.shape[:-1], n_obj + n_bg)
            dw = dw.contiguous()
        _lib.check(_L().na_neus_render_bwd(ptr(sdf), ptr(grad), ptr(dists_obj), ptr(ray_unit), ptr(dist_far), ptr(s_var),
                                           ptr(ctx.anneal_ratio), ptr(rgb_obj), n_obj, ptr(dens_bg), ptr(dists_bg),
                                           ptr(rgb_bg), n_bg, rays, ptr(f32c(d_rgb)), ptr(dw), ptr(d_sdf), ptr(d_grad),
                                           ptr(d_s), ptr(d_rgb_obj), ptr(d_dens), ptr(d_rgb_bg), _lib.current_stream()),
                   "na_neus_render_bwd")
        return d_sdf, d_grad, d_s, d_rgb_obj, d_dens, d_rgb_bg, None, None, None, None, None, None


def neus_render(ray_unit, sdfs, gradients, dists, s_var, rgbs, dist_far=None, progress=1., anneal_end=0.1,
                densities_bg=None, dists_bg=None, rgbs_bg=None, want_weights=True, anneal_ratio=None):
    """One kernel for model.py:132-141: returns (rgb [B,R,3], weights [B,R,No+Nb,1] | None, opacity_all [B,R,1]).
    Object inputs as `neus_alphas` plus rgbs [B,R,No,3]; background: densities [B,R,Nb], dists [B,R,Nb,1], rgbs."""
    anneal_ratio = _anneal_tensor(progress, anneal_end, anneal_ratio, sdfs.device)
    far = None if dist_far is None else dist_far.reshape(sdfs.shape[:-2])
    bg = densities_bg is not None
    rgb, weights, opacity = _NeusRender.apply(
        sdfs[..., 0], gradients, s_var, rgbs, densities_bg if bg else None, rgbs_bg if bg else None,
        dists[..., 0].detach(), ray_unit.detach(), far, dists_bg[..., 0].detach() if bg else None, anneal_ratio,
        want_weights)
    return rgb, (weights[..., None] if want_weights else None), opacity[..., None]
