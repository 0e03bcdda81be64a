"""Autograd wrappers of the element-wise glue kernels (csrc/glue.cu): each replaces a chain of 10-40 few-microsecond
ATen kernels of the reference's per-ray pipeline, forward and backward, by one launch per direction.

    rgb_input(points, ray_unit, gradients, feats, app, sh_levels)   NeuralRGB's input row (modules.py:209-229 + model.py:181)
    tap_combine(taps, sdf, n_taps, eps, training)                   NeuralSDF.compute_gradients' arithmetic (modules.py:144-175)
    sdf_losses(gradients, hessians, outside)                        eikonal / curvature losses (misc.py:74-89)
"""
import numpy as np
import torch

from . import _lib
from ._lib import f32c, ptr

_L = _lib.lib


class _RgbInput(torch.autograd.Function):
    @staticmethod
    def forward(ctx, points, ray_unit, grad, feat, app, samples_per_ray, sh_levels, pad_to=1):
        _lib.require_cuda(points, ray_unit, grad, feat)
        points, ray_unit, grad = f32c(points), f32c(ray_unit), f32c(grad)
        feat = feat if (feat.dtype == torch.float32 and feat.stride(-1) == 1) else f32c(feat)
        rows, n_feat = feat.shape[0], feat.shape[1]
        n_app = 0 if app is None else app.shape[1]
        if app is not None:
            app = f32c(app)
        n_sh = (sh_levels + 1) ** 2
        width = 6 + n_sh + n_feat + n_app
        ld = (width + pad_to - 1) // pad_to * pad_to       # the kernel zero-fills columns width .. ld-1
        out = torch.empty(rows, ld, dtype=torch.float32, device=feat.device)
        _lib.check(_L().na_rgb_input_fwd(ptr(points), ptr(ray_unit), ptr(grad), ptr(feat), feat.stride(0), ptr(app),
                                         0 if app is None else app.stride(0), rows, int(samples_per_ray), int(sh_levels),
                                         n_feat, n_app, ptr(out), ld, _lib.current_stream()), "na_rgb_input_fwd")
        ctx.save_for_backward(grad)
        ctx.cols = (3 + n_sh, 6 + n_sh, n_feat, n_app)
        return out

    @staticmethod
    def backward(ctx, d_out):
        (grad,) = ctx.saved_tensors
        normal_col, feat_col, n_feat, n_app = ctx.cols
        d_out = d_out if (d_out.dtype == torch.float32 and d_out.stride(-1) == 1) else f32c(d_out)
        rows = grad.shape[0]
        d_grad = None
        if ctx.needs_input_grad[2]:
            d_grad = torch.empty_like(grad)
            _lib.check(_L().na_rgb_input_bwd(ptr(grad), ptr(d_out), d_out.stride(0), normal_col, rows, ptr(d_grad),
                                             _lib.current_stream()), "na_rgb_input_bwd")
        d_feat = d_out[:, feat_col:feat_col + n_feat] if ctx.needs_input_grad[3] else None       # a view: no copy
        d_app = d_out[:, feat_col + n_feat:feat_col + n_feat + n_app] if (n_app and ctx.needs_input_grad[4]) else None
        return None, None, d_grad, d_feat, d_app, None, None, None


def rgb_input(points, ray_unit, gradients, feats, app, sh_levels, pad_to=1):
    """points / gradients [B,R,N,3], ray_unit [B,R,3], feats [B,R,N,F], app [B,R,N,A] | None ->
    [B,R,N, 3 + (sh_levels+1)^2 + 3 + F + A]: cat([points, SH(ray_unit), normalize(gradients), feats, app]), with zero
    columns appended up to a multiple of `pad_to` (mlp.linear pads the first layer's weight to match)."""
    lead = feats.shape[:-1]
    n = feats.shape[-2]
    out = _RgbInput.apply(points.reshape(-1, 3), ray_unit.reshape(-1, 3), gradients.reshape(-1, 3),
                          feats.reshape(-1, feats.shape[-1]), None if app is None else app.reshape(-1, app.shape[-1]),
                          n, sh_levels, pad_to)
    return out.view(*lead, out.shape[-1])


class _TapCombine(torch.autograd.Function):
    @staticmethod
    def forward(ctx, taps, sdf, n_taps, den_grad, den_hess, with_hess):
        _lib.require_cuda(taps)
        ctx.shapes = (taps.shape, None if sdf is None else sdf.shape)
        taps = f32c(taps).reshape(n_taps, -1)
        rows = taps.shape[1]
        sdf_c = f32c(sdf).reshape(-1) if with_hess else None
        grad = torch.empty(rows, 3, dtype=torch.float32, device=taps.device)
        hess = torch.empty(rows, 3, dtype=torch.float32, device=taps.device) if with_hess else None
        _lib.check(_L().na_tap_combine_fwd(ptr(taps), ptr(sdf_c), rows, n_taps, den_grad, den_hess, ptr(grad), ptr(hess),
                                           _lib.current_stream()), "na_tap_combine_fwd")
        ctx.args = (n_taps, den_grad, den_hess, with_hess, rows)
        ctx.set_materialize_grads(False)
        if hess is None:
            hess = torch.empty(0, device=taps.device)
            ctx.mark_non_differentiable(hess)
        return grad, hess

    @staticmethod
    def backward(ctx, d_grad, d_hess):
        n_taps, den_grad, den_hess, with_hess, rows = ctx.args
        dev = (d_grad if d_grad is not None else d_hess).device
        d_grad = f32c(d_grad) if d_grad is not None else None
        d_hess = f32c(d_hess) if (d_hess is not None and with_hess) else None
        d_taps = torch.empty(n_taps, rows, dtype=torch.float32, device=dev)
        d_sdf = torch.empty(rows, dtype=torch.float32, device=dev) if (with_hess and ctx.needs_input_grad[1]) else None
        _lib.check(_L().na_tap_combine_bwd(ptr(d_grad), ptr(d_hess), rows, n_taps, den_grad, den_hess, ptr(d_taps),
                                           ptr(d_sdf), _lib.current_stream()), "na_tap_combine_bwd")
        taps_shape, sdf_shape = ctx.shapes
        return d_taps.view(taps_shape), (d_sdf.view(sdf_shape) if d_sdf is not None else None), None, None, None, None


def tap_combine(taps, sdf, n_taps, eps, training):
    """taps [T, ..., 1] (the SDF at x + tap_t), sdf [..., 1] -> (gradient [...,3], hessian [...,3] | None) with the
    reference's arithmetic (modules.py:144-175): eps is the python float the reference divides by."""
    lead = taps.shape[1:-1]
    den_grad = float(np.float32((2 if n_taps == 6 else 4.0) * eps))
    den_hess = float(np.float32(eps ** 2))
    grad, hess = _TapCombine.apply(taps, sdf if training else None, n_taps, den_grad, den_hess, bool(training))
    return grad.view(*lead, 3), (hess.view(*lead, 3) if training else None)


class _SdfLosses(torch.autograd.Function):
    @staticmethod
    def forward(ctx, grad, hess, outside, samples_per_ray):
        _lib.require_cuda(grad)
        ctx.shapes = (grad.shape, None if hess is None else hess.shape)
        grad = f32c(grad).reshape(-1, 3)
        hess = f32c(hess).reshape(-1, 3) if hess is not None else None
        mask = outside.reshape(-1).to(torch.uint8).contiguous() if outside is not None else None
        rows = grad.shape[0]
        out = torch.zeros(2, dtype=torch.float32, device=grad.device)
        _lib.check(_L().na_sdf_losses_fwd(ptr(grad), ptr(hess), ptr(mask), rows, int(samples_per_ray), ptr(out),
                                          _lib.current_stream()), "na_sdf_losses_fwd")
        ctx.save_for_backward(grad, hess, mask)
        ctx.samples = int(samples_per_ray)
        return out

    @staticmethod
    def backward(ctx, d_out):
        grad, hess, mask = ctx.saved_tensors
        rows = grad.shape[0]
        d_out = f32c(d_out)
        d_grad = torch.empty_like(grad)
        d_hess = torch.empty_like(hess) if hess is not None else None
        _lib.check(_L().na_sdf_losses_bwd(ptr(grad), ptr(hess), ptr(mask), rows, ctx.samples, ptr(d_out), ptr(d_grad),
                                          ptr(d_hess), _lib.current_stream()), "na_sdf_losses_bwd")
        g_shape, h_shape = ctx.shapes
        return d_grad.view(g_shape), (d_hess.view(h_shape) if d_hess is not None else None), None, None


def sdf_losses(gradients, hessians, outside):
    """(eikonal_loss, curvature_loss | None) of misc.py:74-89 in one launch.  gradients / hessians [B,R,N,3],
    outside [B,R,1] bool | None."""
    n = gradients.shape[-2]
    out = _SdfLosses.apply(gradients, hessians, outside, n)
    return out[0], (out[1] if hessians is not None else None)
